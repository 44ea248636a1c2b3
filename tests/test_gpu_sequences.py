"""Independent sequences fitted concurrently on one GPU (BASELINE.json configs[4], frame_batch.SequenceStreams): every sequence on
its own stream must reproduce the fit it gives when it runs alone, and the fits must match the oracle."""
import numpy as np
import pytest
import torch

from util import oracle_fit, rel_close, scene_with_target

pytestmark = pytest.mark.gpu


def _field(nnrt, s):
    G = nnrt.geometry
    return G.HierarchicalGraphWarpField.from_graph(s["nodes"], s["edges"], s["arrow_base"], s["node_coverage"], False, 4, 0,
                                        G.WarpNodeCoverageComputationMethod.FIXED_NODE_COVERAGE, 4, s["edge_layer_indices"], s["layer_radii"])


def test_concurrent_sequences_equal_serial_fits(nnrt, oracle):
    from dynamicfuion_python_b200 import frame_batch
    iterations, count, frames = 1, 6, 4  # one iteration: free-running iterations amplify the run-to-run order of the float atomics
    # every sequence tracks its own target (different motion scale), so a cross-talk between workspaces or streams cannot go unnoticed
    scenes = [scene_with_target(oracle, "C2r", 240, 320, translation_scale=0.4 + 0.2 * k) for k in range(count)]
    args = dict(max_iteration_count=iterations, preconditioning_dampening_factor=scenes[0]["lm_factor"], arap_term_weight=scenes[0]["arap_term_weight"])

    def fit(fitter, wf, s):
        eye = torch.eye(3, device=wf.node_rotations.device).repeat(s["nodes"].shape[0], 1, 1)
        wf.node_rotations.copy_(eye); wf.node_translations.zero_()
        fitter.fit_to_image(wf, (s["positions"], s["normals"], s["triangles"]), None, s["reference_points"], s["reference_mask"], s["K"], np.eye(4),
                            (s["height"], s["width"]), anchors=s["anchors"], anchor_weights=s["anchor_weights"])

    serial = []
    for s in scenes:
        fitter, wf = nnrt.alignment.DeformableMeshToImageFitter(**args), _field(nnrt, s)
        fit(fitter, wf, s)
        fitter.check_status()
        serial.append((wf.node_rotations.cpu().numpy().copy(), wf.node_translations.cpu().numpy().copy()))

    pool = frame_batch.SequenceStreams(count)
    resident = [(nnrt.alignment.DeformableMeshToImageFitter(**args), _field(nnrt, s)) for s in scenes]
    pool.fork()
    for frame in range(frames):  # the same frame several times over: the streams stay busy, the last fit is the one compared
        for k, (fitter, wf) in enumerate(resident):
            pool.run(k, fit, fitter, wf, scenes[k])
    pool.join()
    torch.cuda.current_stream().synchronize()
    for k, (fitter, wf) in enumerate(resident):
        fitter.check_status()
        R, t = wf.node_rotations.cpu().numpy(), wf.node_translations.cpu().numpy()
        # float atomics order the sums differently from run to run: tolerance of the node-update parity bar (1e-3 relative)
        assert np.abs(R - serial[k][0]).max() <= 1e-5, k
        ok, worst = rel_close(t, serial[k][1], 1e-3, floor_scale=1e-2)
        assert ok, (k, worst)
    distinct = [np.abs(serial[k][1] - serial[0][1]).max() for k in range(1, count)]
    assert min(distinct) > 1e-5  # the sequences really solved different problems

    expect = oracle_fit(oracle, scenes[-1], iterations)
    ok, worst = rel_close(resident[-1][1].node_translations.cpu().numpy(), expect["translations"], 1e-3, floor_scale=1e-2)
    assert ok, worst


def test_sequence_streams_argument_errors():
    from dynamicfuion_python_b200 import frame_batch
    with pytest.raises(ValueError):
        frame_batch.SequenceStreams(0)
