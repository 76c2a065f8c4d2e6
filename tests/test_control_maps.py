"""Control-map kernels (per-pixel stages): the oracle restatement against fixtures produced by the
reference's own functions (CPU), the kernels against the oracle and the fixtures (GPU).  Integer/byte outputs are
bit-exact; the Gaussian projection involves exp and a max-normalisation, tolerance: alpha 2e-6 abs, uint8 within 1."""
import os

import numpy as np
import pytest
import torch

from oracle import control_maps_restate as cr
from tests.golden.gen_golden_control_maps import gaussians, scene

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
IDX = {"1": 0, "2": 3, "3": 6, "8": 1, "9": 2}


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(G, "control_maps.npz"))


def _scene(gold):
    H, W, N = (int(v) for v in gold["shape"])
    return scene(int(gold["scene_seed"]), H, W, N), (H, W, N)


def test_oracle_composite_and_mask_match_reference(gold):
    (bg_rgb, bg_depth, bg_mask, fg_rgb, fg_depth, fg_mask), (H, W, N) = _scene(gold)
    for i in range(N):
        rgb, depth = cr.composite_by_depth(bg_rgb[i], bg_depth[i], fg_rgb[i], fg_depth[i], fg_mask[i])
        assert np.array_equal(rgb.numpy(), gold["comp_rgb"][i]) and np.array_equal(depth.numpy(), gold["comp_depth"][i])
        m = cr.merge_bg_and_fg_mask(bg_depth[i], fg_depth[i], bg_mask[i], fg_mask[i])
        assert np.array_equal(m.numpy(), gold["merged_mask"][i])


def test_oracle_depth_gray_matches_reference(gold):
    (_, bg_depth, *_), (H, W, N) = _scene(gold)
    for tag in ("norm", "raw"):
        lo, hi = (float(v) for v in gold[f"gray_{tag}_range"])
        for i in range(N):
            assert np.array_equal(cr.depth_to_gray(bg_depth[i], lo, hi).numpy(), gold[f"gray_{tag}"][i])


def test_oracle_gaussian_projection_matches_reference(gold):
    (bg_rgb, *_), (H, W, N) = _scene(gold)
    gp = gaussians(int(gold["gauss_seed"]), N, 3)
    K = torch.from_numpy(gold["K"])
    for f in range(N):
        T = torch.from_numpy(gold["T"][f])
        rgb, alpha = cr.project_frame(list(gp[f].values()), [cr.tab20(IDX[k]) for k in gp[f]], K, T, W, H, 0.003)
        assert np.array_equal(rgb.numpy(), gold["g_rgb"][f])
        assert np.array_equal(alpha.numpy(), gold["g_alpha"][f])
        assert np.array_equal(cr.blend_with_bg(rgb, alpha, bg_rgb[f]).numpy(), gold["g_blend"][f])


def test_host_colour_table_and_cpu_refusal():
    from versecrafter_b200 import _lib, control_maps as cm
    assert cm.TAB20_255 == cr.TAB20_255 and cm.get_object_color("2", IDX, return_float=True) == cr.tab20(3)
    with pytest.raises(_lib.VCError):
        cm.visualize_depth_as_grayscale(torch.zeros(2, 4, 4), 0.5, 2.0)


@pytest.mark.gpu
def test_composite_and_mask_kernels_bit_exact(gold):
    from versecrafter_b200 import control_maps as cm
    (bg_rgb, bg_depth, bg_mask, fg_rgb, fg_depth, fg_mask), _ = _scene(gold)
    c = lambda t: t.cuda()
    rgb, depth, mask = cm.composite_by_depth(c(bg_rgb), c(bg_depth), c(fg_rgb), c(fg_depth), c(fg_mask), bg_mask=c(bg_mask))
    assert np.array_equal(rgb.cpu().numpy(), gold["comp_rgb"]) and np.array_equal(depth.cpu().numpy(), gold["comp_depth"])
    assert np.array_equal(mask.cpu().numpy(), gold["merged_mask"])
    assert torch.equal(cm.merge_bg_and_fg_mask(c(bg_depth), c(fg_depth), c(bg_mask), c(fg_mask)), mask)


@pytest.mark.gpu
def test_depth_gray_kernel_bit_exact(gold):
    from versecrafter_b200 import control_maps as cm
    (_, bg_depth, *_), _ = _scene(gold)
    for tag in ("norm", "raw"):
        lo, hi = (float(v) for v in gold[f"gray_{tag}_range"])
        assert np.array_equal(cm.visualize_depth_as_grayscale(bg_depth.cuda(), lo, hi).cpu().numpy(), gold[f"gray_{tag}"])
    # a full-size frame stack against the oracle (81 x 480 x 832 would take the oracle seconds; 9 frames suffice)
    g = torch.Generator().manual_seed(5)
    d = torch.where(torch.rand(9, 480, 832, generator=g) > 0.1, 0.3 + 20 * torch.rand(9, 480, 832, generator=g), torch.zeros(()))
    got = cm.visualize_depth_as_grayscale(d.cuda(), 0.41, 17.3).cpu()
    assert all(torch.equal(got[i], cr.depth_to_gray(d[i], 0.41, 17.3)) for i in range(9))


@pytest.mark.gpu
def test_gaussian_projection_kernels(gold):
    from versecrafter_b200 import control_maps as cm
    (bg_rgb, *_), (H, W, N) = _scene(gold)
    gp = gaussians(int(gold["gauss_seed"]), N, 3)
    rgb, alpha, blend, masked = cm.project_3d_gaussians_to_2d(gp, IDX, [gold["K"]] * N, list(gold["T"]), (W, H),
                                                              threshold=0.003, background_frames=bg_rgb.cuda())
    assert float((alpha.cpu() - torch.from_numpy(gold["g_alpha"])).abs().max()) <= 2e-6
    for got, want in ((rgb, gold["g_rgb"]), (blend, gold["g_blend"])):
        diff = (got.cpu().int() - torch.from_numpy(want).int()).abs()
        assert int(diff.max()) <= 1 and float((diff > 0).float().mean()) < 0.01
    want_masked = torch.stack([cr.masked_projection(torch.from_numpy(gold["g_rgb"][f]), torch.from_numpy(gold["g_alpha"][f]))
                               for f in range(N)])
    assert int((masked.cpu().int() - want_masked.int()).abs().max()) <= 1
    rgb2, alpha2 = cm.project_3d_gaussians_to_2d(gp, IDX, [gold["K"]] * N, list(gold["T"]), (W, H), threshold=0.003)
    assert torch.equal(rgb2, rgb) and torch.equal(alpha2, alpha)
