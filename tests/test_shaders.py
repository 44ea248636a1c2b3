"""FlatEdgeShader / VertexColorShader over rasteriser fragments (cpp/rendering/kernel/FlatEdgeShaderImpl.h,
cpp/rendering/VertexColorShader.cpp): 8-bit images, bit-exact against the oracle on fragments rasterised from a seeded scene."""
import numpy as np
import pytest

from dynamicfuion_python_b200 import synthetic


def _fragments(oracle, k):
    h, w = 120, 160
    s = synthetic.make_scene("tinyr", h, w)
    fv, mask = oracle.face_vertices_ndc(s["positions"], s["triangles"], s["K"], h, w, 0.0, 10.0)
    idx, depth, bary, dist, _ = oracle.rasterize(fv, mask, h, w, 2.0, k, perspective_correct=True)
    return s, idx, depth, bary, dist


def test_oracle_shaders_against_numpy(oracle):
    s, idx, depth, bary, dist = _fragments(oracle, 2)
    image = oracle.shade_edges_flat(idx, dist, 1.5, (1.0, 0.5, 0.25))
    w2 = np.float32(1.5 / 60.0) ** 2
    d = np.abs(dist[..., 0])
    lit = (idx[..., 0] >= 0) & (d < w2)
    assert lit.any() and not image[~lit].any()
    factor = (w2 - d[lit]) / w2
    assert np.abs(image[lit][:, 0].astype(np.float64) - np.floor(255.0 * factor)).max() <= 1
    rng = np.random.default_rng(4)
    colors = rng.uniform(0, 1, (len(s["positions"]), 3)).astype(np.float32)
    shaded = oracle.shade_vertex_colors(idx, bary, colors[s["triangles"]])
    covered = idx[..., 0] >= 0
    want = (bary[..., 0, :, None] * colors[s["triangles"]][np.maximum(idx[..., 0], 0)]).sum(-2) * 255.0
    assert not shaded[~covered].any()
    inside = covered & (want.min(-1) >= 0) & (want.max(-1) < 255)  # blurred fragments lie outside their face: colours leave [0, 255] and wrap
    assert inside.sum() > 500
    assert np.abs(shaded[inside].astype(np.float64) - np.floor(want[inside])).max() <= 1
    wrapped = np.trunc(want[covered]).astype(np.int64) & 0xFF  # low byte of the truncated value, as the reference's x86 cast
    diff = np.abs(shaded[covered].astype(np.int64) - wrapped)
    assert (np.minimum(diff, 256 - diff) <= 1).all()


@pytest.mark.gpu
@pytest.mark.parametrize("k", [1, 4])
def test_shaders_bit_exact(nnrt, oracle, k):
    s, idx, depth, bary, dist = _fragments(oracle, k)
    shader = nnrt.rendering.FlatEdgeShader(1.5, (1.0, 0.5, 0.25))
    got = shader.shade_meshes(idx, depth, bary, dist)
    assert np.array_equal(got.cpu().numpy(), oracle.shade_edges_flat(idx, dist, 1.5, (1.0, 0.5, 0.25)))
    assert abs(shader.get_ndc_line_width(120, 160) - 1.5 / 60.0) < 1e-9
    rng = np.random.default_rng(4)
    colors = rng.uniform(0, 1, (len(s["positions"]), 3)).astype(np.float32)
    got = nnrt.rendering.VertexColorShader().shade_meshes(idx, depth, bary, dist, [dict(colors=colors, triangles=s["triangles"])])
    assert np.array_equal(got.cpu().numpy(), oracle.shade_vertex_colors(idx, bary, colors[s["triangles"]]))
    with pytest.raises(RuntimeError):
        nnrt.rendering.VertexColorShader().shade_meshes(idx, depth, bary, dist)
    with pytest.raises(RuntimeError):
        nnrt.rendering.FlatEdgeShader(0.0, (1, 1, 1)).shade_meshes(idx, depth, bary, dist)
