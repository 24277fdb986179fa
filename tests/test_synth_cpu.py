"""CPU tier: the synthetic-input generators agree with what the reference would read from disk."""
import numpy as np

from conftest import needs_ref
from pathtracing_b200 import synth


@needs_ref
def test_rgbe_roundtrip_matches_the_reference_hdr_codec():
    """FloatImageTexture texels = Radiance .hdr written and read back by the reference's vendored stb."""
    from oracle import refapi
    lib = refapi.load("strict")
    rng = np.random.default_rng(0)
    wide = (rng.random((64, 48, 3), dtype=np.float32) * np.float32(10) ** rng.integers(-6, 6, (64, 48, 1)).astype(np.float32)).astype(np.float32)
    for img in (synth.hdr_sky(128, 64), wide):
        raw = img if img is wide else img      # hdr_sky is already a fixed point of the codec
        out = np.empty_like(raw)
        assert lib.ref_hdr_roundtrip(raw.ctypes.data, raw.shape[1], raw.shape[0], 3, out.ctypes.data) == 0
        assert np.array_equal(out, synth.rgbe_roundtrip(raw))
    sky = synth.hdr_sky(128, 64)
    assert np.array_equal(synth.rgbe_roundtrip(sky), sky) and sky.max() > 100.0


def test_full_scale_scene_builders_have_the_configured_size():
    d, _ = synth.c3_scene(subdiv=3, n_instances=10)
    assert len(d.instances) == 12 and len(d.blas) == 6
    d, _ = synth.c5_scene(grid=65, n_textures=16, tex_size=8)
    tris = sum(int(m.n_triangles) for m in d.meshes[:16])
    assert tris == 2 * 64 * 64
