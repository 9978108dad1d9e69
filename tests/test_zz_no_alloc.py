"""GPU: bmhrl_rnn_layer_noalloc_f32 -- the recurrent layer of the frozen critic with its workspace in a file-scope device
array (no allocation by caller or library; no counterpart in the reference)
-- against bmhrl_rnn_layer_f32 with a caller workspace, through the C ABI.  Two calls in a row (the shared scratch keeps
the first call's contents), LSTM and GRU, bit for bit (both run the same kernel; the exact-fp32 tests' bound is 1e-6);
shapes beyond the built-in scratch are refused.  Collected last on purpose."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _layer(lib, name, cell, gx, w_hh, b_hh, B, L, H, ws=None):
    out = torch.full((B, L, H), float("nan"), device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    if ws is None:
        lib.call(name, cell, gx.data_ptr(), w_hh.data_ptr(), b_hh.data_ptr(), out.data_ptr(), B, L, H, st)
    else:
        lib.call(name, cell, gx.data_ptr(), w_hh.data_ptr(), b_hh.data_ptr(), out.data_ptr(), ws.data_ptr(), B, L, H, st)
    return out


@pytest.mark.parametrize("cell,NG", [(0, 4), (1, 3)])
def test_rnn_layer_noalloc_equals_workspace_variant(gpu_lib, cell, NG):
    g = torch.Generator().manual_seed(90 + cell)
    B, L, H = 16, 20, 600
    w_hh = (torch.randn(NG * H, H, generator=g) / math.sqrt(H)).cuda()
    b_hh = (torch.randn(NG * H, generator=g) * 0.1).cuda()
    for call in range(2):   # second call: other inputs, the scratch still holds the first call's state
        gx = torch.randn(B, L, NG * H, generator=g).cuda()
        ws = torch.empty(gpu_lib.raw("bmhrl_rnn_ws_floats")(B, H), device="cuda")
        want = _layer(gpu_lib, "bmhrl_rnn_layer_f32", cell, gx, w_hh, b_hh, B, L, H, ws=ws)
        got = _layer(gpu_lib, "bmhrl_rnn_layer_noalloc_f32", cell, gx, w_hh, b_hh, B, L, H)
        torch.cuda.synchronize()
        assert torch.isfinite(got).all(), call
        err = float((got - want).abs().max())
        assert err <= 1e-6, (call, err)
        assert torch.equal(got, want), call
    # a smaller shape right after a larger one (stale contents beyond the new extent), batch not a multiple of 16
    B2, L2 = 3, 5
    gx = torch.randn(B2, L2, NG * H, generator=g).cuda()
    ws = torch.empty(gpu_lib.raw("bmhrl_rnn_ws_floats")(B2, H), device="cuda")
    want = _layer(gpu_lib, "bmhrl_rnn_layer_f32", cell, gx, w_hh, b_hh, B2, L2, H, ws=ws)
    got = _layer(gpu_lib, "bmhrl_rnn_layer_noalloc_f32", cell, gx, w_hh, b_hh, B2, L2, H)
    torch.cuda.synchronize()
    assert torch.equal(got, want)


def test_rnn_layer_noalloc_rejects_shapes_beyond_the_scratch(gpu_lib):
    cap = gpu_lib.raw("bmhrl_rnn_noalloc_max_floats")()
    assert cap == 256 * 1024
    H = 600
    B = cap // H + 1          # bmhrl_rnn_ws_floats(B, H) > cap for sure
    assert gpu_lib.raw("bmhrl_rnn_ws_floats")(B, H) > cap
    x = torch.zeros(8, device="cuda")
    with pytest.raises(RuntimeError, match="built-in scratch"):
        gpu_lib.call("bmhrl_rnn_layer_noalloc_f32", 1, x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), B, 1, H,
                     torch.cuda.current_stream().cuda_stream)
