"""Host-side pieces around the frame: look_at, the gaussian taps, the PNG writer, the oracle's downsample."""

import zlib

import numpy as np

from oracle import nerftex_oracle as O


def test_look_at_is_a_rigid_camera_frame():
    from nerf_tex_b200.dataset import look_at
    for pos in ([3., 0., 0.], [1., -2., 2.5], [0.2, 0.1, 6.]):
        m = look_at(pos, offset=[0.5, 0., -1.])
        R = m[:3, :3]
        np.testing.assert_allclose(R.T @ R, np.eye(3), atol=1e-4)                       # orthonormal up to the eps nudges (:234-236)
        assert abs(np.linalg.det(R) - 1) < 1e-4                                         # right-handed
        np.testing.assert_allclose(m[:3, 2], np.asarray(pos, np.float32) / np.linalg.norm(pos), atol=1e-5 + 1e-6)   # looks at the origin along -z
        assert abs(m[2, 0]) < 1e-4                                                      # right vector stays horizontal (z up)
        np.testing.assert_allclose(m[:3, 3], np.asarray(pos, np.float32) + [0.5, 0., -1.], atol=1e-6)
        np.testing.assert_array_equal(m[3], [0, 0, 0, 1])


def test_gaussian_taps_match_the_2d_kernel():
    from nerf_tex_b200.image import gaussian_taps
    for f in (2, 3, 4, 8):
        t = gaussian_taps(f)
        assert len(t) == int(f * 0.5 * 6) and abs(t.sum() - 1) < 1e-6
        np.testing.assert_allclose(np.outer(t, t), O.gaussian_kernel_2d(len(t), f * 0.5), rtol=1e-5, atol=1e-9)


def test_filtered_downsample_oracle_properties():
    """A constant image stays constant away from the zero-padded border; sizes follow 'SAME' (ceil)."""
    img = np.full((17, 22, 4), 0.7, np.float32)
    out = O.filtered_downsample(img, 3)
    assert out.shape == (6, 8, 4)
    np.testing.assert_allclose(out[2:-2, 2:-2], 0.7, rtol=1e-5)
    assert out[0, 0, 0] < 0.7                                                            # border loses mass to the padding
    lin = np.tile(np.arange(24, dtype=np.float32)[None, :, None], (24, 1, 1))
    out = O.filtered_downsample(lin, 2)
    # 6 taps at x = -2..3 (even sizes are shifted by +0.5, interpolate.py:70), pad 2: output = 2 ox + E[x]
    xs = np.arange(-2, 4)
    wts = np.exp(-0.5 * xs ** 2.0)
    np.testing.assert_allclose(out[5, 3:9, 0], 2 * np.arange(3, 9) + (xs * wts).sum() / wts.sum(), rtol=1e-5)


def test_png_roundtrip_and_structure():
    from nerf_tex_b200.image import decode_png, encode_png
    rng = np.random.default_rng(0)
    for shape in ((5, 7, 4), (16, 16, 3), (3, 2, 1)):
        a = rng.integers(0, 256, shape, dtype=np.uint8)
        data = encode_png(a)
        assert data[:8] == b"\x89PNG\r\n\x1a\n" and data[12:16] == b"IHDR" and data[-8:-4] == b"IEND"
        # every chunk's CRC is valid
        pos = 8
        while pos < len(data):
            n = int.from_bytes(data[pos:pos + 4], "big")
            assert zlib.crc32(data[pos + 4:pos + 8 + n]) == int.from_bytes(data[pos + 8 + n:pos + 12 + n], "big")
            pos += 12 + n
        np.testing.assert_array_equal(decode_png(data), a)


def test_to_uint8():
    x = np.array([-0.2, 0.0, 0.5, 0.999, 1.0, 1.7], np.float32)
    np.testing.assert_array_equal(O.to_uint8(x), [0, 0, 127, 255, 255, 255])
