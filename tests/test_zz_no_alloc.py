"""yh_loss_fwd_bwd_static -- the loss forward + backward for callers that cannot allocate
(no counterpart in the reference).  Its scratch is a 128 KiB file-scope device array inside
libyolohead.so.  Called here straight through the C ABI (ctypes) and held to the existing entry point
yh_loss_fwd_bwd on the same inputs, on calls in a row that share the array.

This file is named so that pytest collects it LAST."""
import ctypes as C

import numpy as np
import pytest
import torch

from yolo_v1_b200 import _lib, synth

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _existing(pred, tgt, S, B, Cn, want_grad=True):
    """The existing entry point, with a caller-supplied (torch) workspace."""
    lib = _lib.lib()
    n = pred.shape[0]
    ws = torch.zeros(int(lib.yh_loss_workspace_bytes(n, S, B, Cn)) + 256, dtype=torch.uint8, device=pred.device)
    pad = (-ws.data_ptr()) % 256
    parts = torch.empty(5, dtype=torch.float32, device=pred.device)
    grad = torch.empty_like(pred) if want_grad else None
    rc = lib.yh_loss_fwd_bwd(pred.data_ptr(), _lib.dtype_code(pred), tgt.data_ptr(), n, S, B, Cn, 5.0, 0.5,
                             parts.data_ptr(), _lib.ptr(grad), ws.data_ptr() + pad, ws.numel() - pad,
                             _lib.stream_ptr(pred.device))
    _lib.check(rc, "yh_loss_fwd_bwd")
    return parts, grad


def _static(pred, tgt, S, B, Cn, want_grad=True):
    """The new entry point: inputs and outputs only, no workspace argument."""
    lib = _lib.lib()
    parts = torch.empty(5, dtype=torch.float32, device=pred.device)
    grad = torch.empty_like(pred) if want_grad else None
    rc = lib.yh_loss_fwd_bwd_static(pred.data_ptr(), _lib.dtype_code(pred), tgt.data_ptr(), pred.shape[0], S, B, Cn,
                                    5.0, 0.5, parts.data_ptr(), _lib.ptr(grad), _lib.stream_ptr(pred.device))
    return rc, parts, grad


def _batch(n, seed, S=7, B=2, Cn=20, dtype=torch.float32):
    tgt = synth.make_targets(n, seed=seed, S=S, B=B, C=Cn, p_obj=0.1, both_slots=0.2 if B > 1 else 0.0)
    pred = synth.make_preds(n, "uniform", seed=seed + 1, S=S, B=B, C=Cn)
    return pred.to(DEV).to(dtype).contiguous(), tgt.to(DEV).contiguous()


def test_static_scratch_loss_matches_the_existing_entry_point_on_calls_in_a_row():
    lib = _lib.lib()
    assert lib.yh_loss_static_scratch_bytes() == 128 * 1024
    # calls in a row on one stream, sharing the array: a large batch, then a small one (stale partials of the first
    # call sit behind it), then the first again; then half precision and another head shape
    calls = [(_batch(6000, 1), (7, 2, 20)), (_batch(37, 3), (7, 2, 20)), (_batch(6000, 1), (7, 2, 20)),
             (_batch(512, 5, dtype=torch.bfloat16), (7, 2, 20)), (_batch(300, 7, S=4, B=3, Cn=5), (4, 3, 5))]
    for i, ((pred, tgt), (S, B, Cn)) in enumerate(calls):
        if i == 2:          # by now every kernel is loaded: the device's free memory must not move across a call
            torch.cuda.synchronize()
            free_before, _ = torch.cuda.mem_get_info()
            parts = torch.empty(5, dtype=torch.float32, device=DEV)
            grad = torch.empty_like(pred)
            torch.cuda.synchronize()
            free_before, _ = torch.cuda.mem_get_info()
            rc = lib.yh_loss_fwd_bwd_static(pred.data_ptr(), _lib.dtype_code(pred), tgt.data_ptr(), pred.shape[0], S, B, Cn,
                                            5.0, 0.5, parts.data_ptr(), grad.data_ptr(), _lib.stream_ptr(pred.device))
            torch.cuda.synchronize()
            free_after, _ = torch.cuda.mem_get_info()
            assert rc == 0 and free_after == free_before, "the call changed the device's free memory"
        else:
            rc, parts, grad = _static(pred, tgt, S, B, Cn)
            assert rc == 0, lib.yh_last_error()
        want_parts, want_grad = _existing(pred, tgt, S, B, Cn)
        # the same kernel on the same inputs: in practice bit-identical; held to the suite's loss tolerance
        np.testing.assert_allclose(parts.cpu().numpy(), want_parts.cpu().numpy(), rtol=1e-5, atol=1e-9, err_msg=f"call {i}")
        np.testing.assert_allclose(grad.float().cpu().numpy(), want_grad.float().cpu().numpy(), rtol=1e-5, atol=1e-7,
                                   err_msg=f"call {i}")
    # forward only (no gradient output)
    (pred, tgt), (S, B, Cn) = calls[1]
    rc, parts, grad = _static(pred, tgt, S, B, Cn, want_grad=False)
    assert rc == 0 and grad is None
    np.testing.assert_allclose(parts.cpu().numpy(), _existing(pred, tgt, S, B, Cn, False)[0].cpu().numpy(), rtol=1e-5)


def test_static_scratch_loss_rejects_what_does_not_fit():
    lib = _lib.lib()
    pred, tgt = _batch(10679, 11)            # one image more than the array holds at S=7
    assert lib.yh_loss_workspace_bytes(10678, 7, 2, 20) <= 128 * 1024 < lib.yh_loss_workspace_bytes(10679, 7, 2, 20)
    rc, _, _ = _static(pred, tgt, 7, 2, 20)
    assert rc == -3 and b"10,678" in lib.yh_last_error()          # YH_ERR_WORKSPACE, with the limit in the message
    rc, parts, _ = _static(pred[:10678], tgt[:10678], 7, 2, 20)  # the largest batch that fits
    assert rc == 0
    np.testing.assert_allclose(parts.cpu().numpy(), _existing(pred[:10678], tgt[:10678], 7, 2, 20)[0].cpu().numpy(), rtol=1e-5)
    assert _static(pred[:4], tgt[:4], 7, 9, 20)[0] == -2          # a shape outside the build
    assert lib.yh_loss_fwd_bwd_static(None, 0, None, 4, 7, 2, 20, 5.0, 0.5, None, None, None) == -1
