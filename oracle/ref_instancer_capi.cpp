// TEST INFRASTRUCTURE ONLY.  C entry points around the REFERENCE's own instancer (instancer/src/instancer.cpp, compiled
// from where it lies under /root/reference; nothing of it is copied here), bound from oracle/ref_instancer.py with ctypes.
// The reference's Cython binding (instancer/instancer.pyx) includes the .cpp the same way.  Embree is replaced by the
// brute-force stand-in in oracle/embree_shim/ (see there for what that does and does not preserve).
#include "src/instancer.cpp"

#include <string>
#include <vector>

extern "C" {

void* ref_instancer_new(const float* b_0, const float* b_1, int cast_shadow_rays, const char* const* textures, int n_textures,
                        int min_shadow_samples, int n_shadow_samples, int min_texture_samples, int n_texture_samples,
                        float jitter_amount, int instance_sample_method, int use_mean_distance, unsigned int seed) {
    std::vector<std::string> tex;
    for (int i = 0; i < n_textures; ++i) tex.push_back(textures[i]);
    return new C_Instancer(std::vector<float>(b_0, b_0 + 3), std::vector<float>(b_1, b_1 + 3), cast_shadow_rays != 0, tex,
                           min_shadow_samples, n_shadow_samples, min_texture_samples, n_texture_samples, jitter_amount,
                           (uint8_t)instance_sample_method, use_mean_distance != 0, seed);
}

void ref_instancer_delete(void* p) { delete (C_Instancer*)p; }

void ref_instancer_add_instance(void* p, const float* m16) {           // row-major 4x4, instancer.pyx: sum(transformation, [])
    std::vector<float> t(m16, m16 + 16);
    ((C_Instancer*)p)->AddInstance(t);
}

int ref_instancer_distribute(void* p, const char* mesh_path, float scale, const char* patch_origins_path) {
    try {
        ((C_Instancer*)p)->DistributeInstancesOnMesh(mesh_path, scale, patch_origins_path);
    } catch (...) {
        return 1;
    }
    return 0;
}

int ref_instancer_add_mesh(void* p, const char* mesh_path, const char* texture_path) {
    try {
        ((C_Instancer*)p)->AddMesh(mesh_path, texture_path);
    } catch (...) {
        return 1;
    }
    return 0;
}

void ref_instancer_commit(void* p) { ((C_Instancer*)p)->CommitScene(); }

unsigned int ref_instancer_count(void* p) { return ((C_Instancer*)p)->GetNumberOfInstances(); }

void ref_instancer_export(void* p, const char* path) { ((C_Instancer*)p)->ExportTransformations(path); }

void ref_instancer_get_model_input(void* p, float* rays_o, float* rays_d, float* t, float* dists, float* pts, float* color,
                                   float* density, float* density_weight, int* instance_id, bool* hit, unsigned int n_rays,
                                   unsigned int n_pts, float step_size, float* parameters) {
    ((C_Instancer*)p)->GetModelInput(rays_o, rays_d, t, dists, pts, color, density, density_weight, instance_id, hit, n_rays, n_pts,
                                     step_size, parameters);
}

}  // extern "C"
