"""TEST INFRASTRUCTURE ONLY: `instancer.instancer.Instancer` for the reference's network/renderer.py when it runs over
the TensorFlow stand-in — the reference's own C++ instancer (oracle/ref_instancer.py) behind the interface of the
reference's Cython binding (instancer/instancer.pyx:6-54): same constructor, get_model_input returns the ten tensors."""

import numpy as np
import tensorflow as tf

from oracle import ref_instancer as RI


class Instancer:
    def __init__(self, b_0, b_1, cast_shadow_rays=False, textures=[], transformations=[], mesh_path=None, patch_scale=1.,
                 patch_origins_path='', min_shadow_samples=4, n_shadow_samples=512, min_texture_samples=4,
                 n_texture_samples=512, jitter_amount=0, instance_sampling_method='random', use_mean_distance=False,
                 auxiliary_meshes=[], transformation_export_path=None):
        self.ref = RI.RefInstancer(list(b_0), list(b_1), cast_shadow_rays, list(textures), list(transformations), mesh_path,
                                   patch_scale, patch_origins_path, min_shadow_samples, n_shadow_samples, min_texture_samples,
                                   n_texture_samples, jitter_amount, instance_sampling_method, use_mean_distance,
                                   list(auxiliary_meshes), transformation_export_path)
        self.last = None

    def n_instances(self):
        return self.ref.n_instances()

    def get_model_input(self, rays_o, rays_d, parameters, n_samples, step_size):
        r = self.ref.get_model_input(rays_o.numpy(), rays_d.numpy(), np.asarray(parameters.numpy() if hasattr(parameters, "numpy") else parameters),
                                     n_samples, step_size)
        self.last = r                                                # what the C++ returned, for the comparing test
        c = tf.constant
        return (c(r["rays_d"]), c(r["pts"]), c(r["t"]), c(r["dist"]), c(r["color"]), c(r["density"]), c(r["density_weight"]),
                c(r["instance_id"]), tf.where(r["hit"]), c(r["parameters"]))
