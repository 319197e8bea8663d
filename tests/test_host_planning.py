"""CPU tests of host-side planning code: the kernel-selection predicates of the native library (pure host functions,
callable without a GPU) and the stride inspection behind the channel-slice gather."""
import ctypes
import os

import pytest
import torch

from mpi_pytorch_b200.build import lib_path


@pytest.fixture(scope="module")
def lib():
    path = lib_path()
    if not os.path.exists(path):
        pytest.skip("native library not built")
    L = ctypes.CDLL(path)
    L.mpb_tc_halo_efficiency.restype = ctypes.c_float
    return L


def test_resnet_layers_select_the_specialised_kernels(lib):
    i = ctypes.c_int
    # persistent halo fprop/dgrad: weights of the CTA's output-channel block must stay resident in shared memory
    assert lib.mpb_tc_halo_persistent_ok(i(64), i(64), i(3), i(3), i(56), i(56)) == 1
    assert lib.mpb_tc_halo_persistent_ok(i(128), i(128), i(3), i(3), i(28), i(28)) == 1
    assert lib.mpb_tc_halo_persistent_ok(i(256), i(256), i(3), i(3), i(14), i(14)) == 0     # 9 x 256 x 64 x 2 B does not fit
    assert lib.mpb_tc_halo_persistent_ok(i(64), i(64), i(4), i(1), i(112), i(112)) == 1      # space-to-depth stem
    # halo wgrad: C == 64 (tap pairs) or C % 128 == 0 (one filter row per CTA), K % 64 == 0, <= 40 column groups
    assert lib.mpb_tc_wgrad_halo_ok(i(64), i(64), i(3), i(3), i(56), i(56)) == 1
    assert lib.mpb_tc_wgrad_halo_ok(i(128), i(128), i(3), i(3), i(28), i(28)) == 1
    assert lib.mpb_tc_wgrad_halo_ok(i(256), i(256), i(3), i(3), i(14), i(14)) == 1
    assert lib.mpb_tc_wgrad_halo_ok(i(512), i(512), i(3), i(3), i(7), i(7)) == 0
    assert lib.mpb_tc_wgrad_halo_ok(i(96), i(64), i(3), i(3), i(12), i(12)) == 0
    assert lib.mpb_tc_wgrad_halo_ok(i(64), i(48), i(3), i(3), i(12), i(12)) == 0
    assert lib.mpb_tc_wgrad_halo_ok(i(64), i(64), i(1), i(1), i(56), i(56)) == 0


def test_halo_tile_geometry_fills_the_128_row_mma(lib):
    i = ctypes.c_int
    for (ho, wo, r, s) in ((56, 56, 3, 3), (28, 28, 3, 3), (112, 112, 4, 1), (14, 14, 3, 3), (7, 7, 3, 3), (224, 224, 3, 3)):
        eff = float(lib.mpb_tc_halo_efficiency(i(ho), i(wo), i(r), i(s)))
        assert 0.0 < eff <= 1.0
    assert float(lib.mpb_tc_halo_efficiency(i(56), i(56), i(3), i(3))) >= 0.85
    assert float(lib.mpb_tc_halo_efficiency(i(112), i(112), i(4), i(1))) >= 0.9


def _row_view_fn():
    """ops/cuda_impl.py binds the CUDA library at import; pull the pure stride-inspection helper out of its source."""
    import mpi_pytorch_b200.ops as ops_pkg
    src = open(os.path.join(os.path.dirname(ops_pkg.__file__), "cuda_impl.py")).read()
    start = src.index("def _row_view(")
    end = src.index("def _copy_rows(")
    ns = {"torch": torch}
    exec(src[start:end], ns)
    return ns["_row_view"]


def test_row_view_recognises_channel_slices_of_nhwc_tensors():
    row_view = _row_view_fn()
    big = torch.randn(2, 160, 5, 6).contiguous(memory_format=torch.channels_last)
    assert row_view(big) == (2 * 5 * 6, 160, 160)
    assert row_view(big[:, 32:96]) == (60, 64, 160)                      # channel slice: pitch of the parent
    assert row_view(big[:, :, 1:4]) is None                              # a spatial crop is not a row view
    assert row_view(torch.randn(2, 160, 5, 6)) is None                   # NCHW-contiguous
    assert row_view(torch.randn(4, 8)) is None


def test_measured_algo_table_lookup():
    """parallel/algo_table.json (written by scripts/peer_sweep.py) drives the per-bucket algorithm / grid choice."""
    from mpi_pytorch_b200.parallel.peer import load_algo_table, table_choice
    table = load_algo_table()
    assert set(table) >= {2, 4, 8}
    for w, rows in table.items():
        assert rows == sorted(rows, key=lambda r: r[0]) and all(a in ("oneshot", "twoshot", "nvls") and 1 <= g <= 64
                                                                for _, a, g in rows)
    assert table_choice(table, 8, 32 << 20, True) == ("nvls", 16)          # the measured optimum at W = 8
    assert table_choice(table, 8, 32 << 20, False)[0] == "twoshot"         # no multicast on this host: same grid, peer loads
    assert table_choice(table, 8, 4 << 30, True) is not None               # beyond the last row: the last row applies
    assert table_choice(table, 3, 1 << 20, True) is None                   # unmeasured world size: fixed crossover
    assert table_choice({}, 8, 1 << 20, True) is None


def test_conv_relu_pairs_are_fused_without_changing_the_model():
    """run_stack fuses Conv->ReLU pairs of the VGG / AlexNet feature stacks; outputs and parameter names are unchanged."""
    import torch
    from mpi_pytorch_b200.models import initialize_model
    for name in ("vgg16", "alexnet"):
        torch.manual_seed(0)
        m, _ = initialize_model(name, 7, False, use_pretrained=False)
        m.eval()
        x = torch.randn(2, 3, 64, 64)
        with torch.no_grad():
            want = m.classifier(torch.flatten(m.avgpool(m.features(x)), 1))     # the plain nn.Sequential walk
            got = m(x)
        assert torch.allclose(got, want, atol=1e-5)
        assert any(k.startswith("features.0.weight") for k in m.state_dict())
