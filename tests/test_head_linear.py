"""The head's last Linear fused with the loss (yh_head_linear_loss) against fixtures the UNMODIFIED
reference produced (tests/golden/make_head_linear_golden.py: nn.Linear(4096, 1470) + view + YOLOLoss + backward, fp32 on
CPU, operands exact in bf16 and fp16), against this package's own loss kernel on the same predictions (bit for bit), and
against torch's Linear on the GPU."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "head_linear.npz")
KEYS = ("total", "coord", "conf_obj", "conf_noobj", "class")
# fp32 accumulation of exact products in another order than the reference's CPU GEMM: ~1e-6 relative on a prediction;
# the loss sums 49 * M cells of them
RTOL_FP32 = 1e-5


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def _inputs(gold, name, dtype):
    from yolo_v1_b200 import synth

    M, seed = int(gold[f"{name}/M"]), int(gold[f"{name}/seed"])
    x, w, b, tgt = synth.head_linear_inputs(M, seed)
    dev = torch.device("cuda")
    return (torch.from_numpy(x).to(dev, dtype), torch.from_numpy(w).to(dev, dtype), torch.from_numpy(b).to(dev),
            torch.from_numpy(tgt).to(dev))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("name", ["m64", "m200", "m333"])
def test_fused_matches_reference_fixture(gold, name, dtype):
    """Loss dict, d loss / d prediction and the prediction itself against the reference's fp32 run (no rounding of the
    prediction: round_pred=False, fp32 gradient)."""
    from yolo_v1_b200.head_linear import linear_loss

    x, w, b, tgt = _inputs(gold, name, dtype)
    parts, grad, pred = linear_loss(x, w, b, tgt, round_pred=False, grad_dtype=torch.float32, want_pred=True)
    np.testing.assert_allclose(parts.cpu().numpy().astype(np.float64), gold[f"{name}/parts"], rtol=RTOL_FP32)
    np.testing.assert_allclose(pred[:4].cpu().numpy(), gold[f"{name}/pred_head"], rtol=1e-5, atol=2e-6)
    M = x.shape[0]
    g = grad.view(M, 7, 7, 30)
    np.testing.assert_allclose(g[:4].cpu().numpy(), gold[f"{name}/dpred_head"], rtol=2e-4, atol=1e-7)
    assert float(g.double().abs().sum()) == pytest.approx(float(gold[f"{name}/dpred_abs_sum"]), rel=1e-5)
    np.testing.assert_allclose(g.reshape(M, 1470).double().sum(0).cpu().numpy(), gold[f"{name}/db"], rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("name", ["m64", "m333"])
def test_module_backward_matches_reference_fixture(gold, name):
    """FusedHeadLoss over an nn.Linear with fp32 master weights: loss.backward() gives the reference's activation, weight
    and bias gradients (fp32 backward GEMMs; the operands are exact in bf16, so the cast loses nothing)."""
    from yolo_v1_b200 import FusedHeadLoss, YOLOLoss

    x, w, b, tgt = _inputs(gold, name, torch.float32)
    lin = torch.nn.Linear(4096, 1470).cuda()
    with torch.no_grad():
        lin.weight.copy_(w)
        lin.bias.copy_(b)
    fused = FusedHeadLoss(lin, YOLOLoss(), operand_dtype=torch.bfloat16, round_pred=False, backward_dtype=torch.float32)
    xr = x.clone().requires_grad_(True)
    loss, d = fused(xr, tgt)
    loss.backward()
    np.testing.assert_allclose([d[k] for k in KEYS], gold[f"{name}/parts"], rtol=RTOL_FP32)
    assert float(loss.detach()) == pytest.approx(float(gold[f"{name}/parts"][0]), rel=RTOL_FP32)
    np.testing.assert_allclose(lin.bias.grad.cpu().numpy(), gold[f"{name}/db"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(xr.grad[:8, :256].cpu().numpy(), gold[f"{name}/dx_head"], rtol=1e-3, atol=1e-7)
    np.testing.assert_allclose(lin.weight.grad[:64, :128].cpu().numpy(), gold[f"{name}/dw_head"], rtol=1e-3, atol=1e-7)
    assert float(xr.grad.double().abs().sum()) == pytest.approx(float(gold[f"{name}/dx_abs_sum"]), rel=1e-4)
    assert float(lin.weight.grad.double().abs().sum()) == pytest.approx(float(gold[f"{name}/dw_abs_sum"]), rel=1e-4)


@pytest.mark.parametrize("name", ["m64", "m200", "m333"])
def test_rounded_prediction_matches_reference_on_rounded_prediction(gold, name):
    """round_pred=True (the default): the loss of the bf16-rounded prediction, as the reference's criterion computes it
    on that prediction widened to fp32; and within the AMP tolerance (2e-2 for bf16, tests/test_gpu_parity.py) of the
    reference's own autocast run, whose criterion also rounds intermediate values to bf16."""
    from yolo_v1_b200.head_linear import linear_loss

    x, w, b, tgt = _inputs(gold, name, torch.bfloat16)
    parts = linear_loss(x, w, b, tgt, round_pred=True)[0].cpu().numpy().astype(np.float64)
    # a prediction that lands within fp32 summation noise of a bf16 rounding boundary may round the other way
    np.testing.assert_allclose(parts, gold[f"{name}/parts_rounded_bf16"], rtol=2e-4)
    np.testing.assert_allclose(parts, gold[f"{name}/parts_autocast_bf16"], rtol=2e-2)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("M", [1, 127, 128, 129, 1000, 4097, 20001])
def test_loss_is_the_loss_kernel_on_the_same_predictions(M, dtype):
    """The epilogue's loss arithmetic is yh_loss_fwd_bwd's: same parts, same gradient bits, on the predictions the kernel
    itself reports; the plain layer (targets=None) reports the same predictions; they match torch's fp32 Linear."""
    from yolo_v1_b200 import head_linear, synth
    from yolo_v1_b200.head_linear import linear_loss
    from yolo_v1_b200.loss import loss_fwd_bwd

    g = torch.Generator(device="cuda").manual_seed(M)
    x = (torch.randn(M, 4096, device="cuda", generator=g) * 0.5).to(dtype)
    w = (torch.randn(1470, 4096, device="cuda", generator=g) * 0.02).to(dtype)
    b = torch.rand(1470, device="cuda", generator=g) * 0.6
    tgt = synth.make_targets(M, seed=M, p_obj=0.2, size_mix=True).cuda()
    for rp in (False, True):
        parts, grad, pred = linear_loss(x, w, b, tgt, round_pred=rp, grad_dtype=torch.float32, want_pred=True)
        want_parts, want_grad = loss_fwd_bwd(pred, tgt, 7, 2, 20, 5.0, 0.5, True)
        # (the loss kernel adds a warp's cells in fp32 before it goes to fp64; this one adds every cell in fp64)
        np.testing.assert_allclose(parts.cpu().numpy(), want_parts.cpu().numpy(), rtol=2e-6)
        assert torch.equal(grad.view_as(want_grad), want_grad)
        assert torch.equal(head_linear(x, w, b, round_pred=rp), pred)
        ref = torch.nn.functional.linear(x.float(), w.float(), b).view(M, 7, 7, 30)
        if rp:
            ref = ref.to(dtype).float()
            bad = (pred != ref).float().mean().item()
            # fp32 summation noise (~5e-6) against the spacing of the narrow type: one unit in the last place of a few
            assert bad < (5e-2 if dtype is torch.float16 else 5e-3) and (pred - ref).abs().max().item() <= 0.04
        else:
            assert (pred - ref).abs().max().item() < 1e-4
    # the narrow gradient is the fp32 one times the scale, rounded once
    scale = 1024.0
    _, g16, _ = linear_loss(x, w, b, tgt, round_pred=False, grad_scale=scale)
    _, g32, _ = linear_loss(x, w, b, tgt, round_pred=False, grad_dtype=torch.float32)
    assert g16.dtype is dtype and torch.equal(g16, (g32 * scale).to(dtype))


def test_module_under_grad_scaler_and_against_two_calls():
    """What a training step does: fp32 master weights, fp16 operands, scaler.scale(loss).backward().  Against the same
    step through torch's autocast Linear + this package's YOLOLoss."""
    from yolo_v1_b200 import FusedHeadLoss, YOLOLoss, synth

    M = 512
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(M, 4096, device="cuda", generator=g) * 0.1   # predicted w, h stay away from 0, where sqrt' amplifies GEMM summation noise
    tgt = synth.make_targets(M, seed=3, p_obj=0.1, size_mix=True).cuda()
    lin = torch.nn.Linear(4096, 1470).cuda()
    with torch.no_grad():
        lin.bias.add_(0.4)
    crit = YOLOLoss()
    scale = 4096.0
    fused = FusedHeadLoss(lin, crit, operand_dtype=torch.float16)
    xr = x.clone().requires_grad_(True)
    loss, d = fused(xr, tgt, loss_scale=scale)
    (loss * scale).backward()
    got = (xr.grad.clone(), lin.weight.grad.clone(), lin.bias.grad.clone())
    lin.zero_grad()
    xr2 = x.clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.float16):
        pred = lin(xr2).view(-1, 7, 7, 30)
    loss2, d2 = crit(pred, tgt)
    (loss2 * scale).backward()
    assert float(loss.detach()) == pytest.approx(float(loss2.detach()), rel=2e-3)
    assert all(d[k] == pytest.approx(d2[k], rel=2e-3, abs=1e-6) for k in KEYS)
    for a, b_ in zip(got, (xr2.grad, lin.weight.grad, lin.bias.grad)):
        assert a.dtype is torch.float32 and torch.isfinite(a).all()
        rel = (a - b_).norm() / b_.norm()
        assert float(rel) < 5e-2, float(rel)   # both round the prediction to fp16; a few round the other way


def test_unsupported_shapes_and_dtypes_raise():
    from yolo_v1_b200 import FusedHeadLoss, YOLOLoss, _lib
    from yolo_v1_b200.head_linear import linear_loss

    x = torch.zeros(4, 4096, device="cuda", dtype=torch.bfloat16)
    w = torch.zeros(1470, 4096, device="cuda", dtype=torch.bfloat16)
    tgt = torch.zeros(4, 7, 7, 30, device="cuda")
    with pytest.raises(NotImplementedError):
        linear_loss(x.double(), w.double(), None, tgt)
    with pytest.raises(NotImplementedError):
        linear_loss(x.float(), w, None, tgt)                      # operands of two types
    with pytest.raises(NotImplementedError):
        linear_loss(x[:, :2048].contiguous(), w[:, :2048].contiguous(), None, tgt)
    with pytest.raises(_lib.YoloHeadError):
        linear_loss(x.cpu(), w, None, tgt)
    with pytest.raises(_lib.YoloHeadError):
        linear_loss(x, w, None, tgt[:3])
    with pytest.raises(NotImplementedError):
        FusedHeadLoss(torch.nn.Linear(4096, 7 * 7 * 13), YOLOLoss(C=3))
    parts, grad, pred = linear_loss(x[:0], w, None, tgt[:0])     # empty batch: nothing launched
    assert grad.shape == (0, 1470)


def test_capturable_in_a_cuda_graph():
    """The call only enqueues work (tensor maps travel as kernel arguments): captured once, replayed on new data."""
    from yolo_v1_b200 import synth
    from yolo_v1_b200.head_linear import linear_loss

    M = 700
    g = torch.Generator(device="cuda").manual_seed(11)
    x = (torch.randn(M, 4096, device="cuda", generator=g) * 0.1).to(torch.bfloat16)
    w = (torch.randn(1470, 4096, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
    b = torch.rand(1470, device="cuda", generator=g) * 0.6
    tgt = synth.make_targets(M, seed=2, p_obj=0.1, size_mix=True).cuda()
    linear_loss(x, w, b, tgt)                       # workspace exists before the capture
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        parts, grad, _ = linear_loss(x, w, b, tgt)
    for seed in (21, 22):
        x.copy_((torch.randn(M, 4096, device="cuda", generator=g.manual_seed(seed)) * 0.1).to(torch.bfloat16))
        graph.replay()
        want_parts, want_grad, _ = linear_loss(x, w, b, tgt)
        assert torch.equal(parts, want_parts) and torch.equal(grad, want_grad)


def test_every_visible_device_in_one_process():
    """Per-device state (function attributes, workspaces, tensor maps): the same call on every GPU this process sees."""
    from yolo_v1_b200 import synth
    from yolo_v1_b200.head_linear import linear_loss

    M = 300
    x, w, b, tgt = synth.head_linear_inputs(M, 9)
    outs = []
    for d in range(torch.cuda.device_count()):
        dev = torch.device("cuda", d)
        args = (torch.from_numpy(x).to(dev, torch.bfloat16), torch.from_numpy(w).to(dev, torch.bfloat16),
                torch.from_numpy(b).to(dev), torch.from_numpy(tgt).to(dev))
        parts, grad, pred = linear_loss(*args, want_pred=True)
        assert parts.device == dev
        outs.append((parts.cpu(), grad.cpu(), pred.cpu()))
    for o in outs[1:]:
        assert all(torch.equal(a, b_) for a, b_ in zip(o, outs[0]))


# ----------------------------------------------------------------------------- fp32 operands (the reference's default step)
@pytest.mark.parametrize("precision", ["3xtf32", "tf32"])
@pytest.mark.parametrize("name", ["m64", "m333"])
def test_fp32_operands_match_reference_fixture(gold, name, precision):
    """The fixtures' operands have 8 significant bits: exact in TF32 too, so both precisions must reproduce the reference."""
    from yolo_v1_b200.head_linear import linear_loss

    x, w, b, tgt = _inputs(gold, name, torch.float32)
    parts, grad, pred = linear_loss(x, w, b, tgt, want_pred=True, precision=precision)
    assert grad.dtype is torch.float32
    np.testing.assert_allclose(parts.cpu().numpy().astype(np.float64), gold[f"{name}/parts"], rtol=RTOL_FP32)
    np.testing.assert_allclose(pred[:4].cpu().numpy(), gold[f"{name}/pred_head"], rtol=1e-5, atol=2e-6)
    M = x.shape[0]
    np.testing.assert_allclose(grad.view(M, 7, 7, 30)[:4].cpu().numpy(), gold[f"{name}/dpred_head"], rtol=2e-4, atol=1e-7)


@pytest.mark.parametrize("M", [300, 4096])
def test_3xtf32_has_fp32_accuracy_on_arbitrary_operands(M):
    """Random fp32 operands (24 significant bits) against an fp64 Linear: 3xTF32 keeps the prediction within 2e-5 absolute
    (torch's own fp32 Linear: 2e-6; what is left is the tensor core's truncating accumulate over 512 steps) and the loss
    within 1e-5 relative -- north_star's tolerance; one TF32 pass is ~5e-4 / 4e-4, what torch's allow_tf32 gives."""
    from yolo_v1_b200 import synth
    from yolo_v1_b200.head_linear import linear_loss
    from yolo_v1_b200.loss import loss_fwd_bwd

    g = torch.Generator(device="cuda").manual_seed(M)
    x = torch.randn(M, 4096, device="cuda", generator=g) * 0.1
    w = torch.randn(1470, 4096, device="cuda", generator=g) * 0.02
    b = torch.rand(1470, device="cuda", generator=g) * 0.6
    tgt = synth.make_targets(M, seed=M, p_obj=0.2, size_mix=True).cuda()
    ref = (x.double() @ w.double().T + b.double()).view(M, 7, 7, 30)
    want_parts = loss_fwd_bwd(ref.float(), tgt, 7, 2, 20, 5.0, 0.5, False)[0].cpu().numpy()
    parts, grad, pred = linear_loss(x, w, b, tgt, want_pred=True)                 # precision="3xtf32" is the default
    assert (pred.double() - ref).abs().max().item() < 2e-5
    np.testing.assert_allclose(parts.cpu().numpy(), want_parts, rtol=1e-5)
    # the loss arithmetic is the loss kernel's on the predictions the kernel reports
    same_parts, same_grad = loss_fwd_bwd(pred, tgt, 7, 2, 20, 5.0, 0.5, True)
    assert torch.equal(grad.view_as(same_grad), same_grad)
    np.testing.assert_allclose(parts.cpu().numpy(), same_parts.cpu().numpy(), rtol=2e-6)
    parts1, _, pred1 = linear_loss(x, w, b, tgt, want_pred=True, precision="tf32")
    err1 = (pred1.double() - ref).abs().max().item()
    assert 5e-5 < err1 < 2e-3
    np.testing.assert_allclose(parts1.cpu().numpy(), want_parts, rtol=2e-3)


def test_module_with_fp32_operands_matches_reference_fixture(gold):
    """FusedHeadLoss(operand_dtype=torch.float32): the reference's non-AMP step, gradients for activations, weight, bias."""
    from yolo_v1_b200 import FusedHeadLoss, YOLOLoss, _lib
    from yolo_v1_b200.head_linear import linear_loss

    name = "m200"
    x, w, b, tgt = _inputs(gold, name, torch.float32)
    lin = torch.nn.Linear(4096, 1470).cuda()
    with torch.no_grad():
        lin.weight.copy_(w)
        lin.bias.copy_(b)
    fused = FusedHeadLoss(lin, YOLOLoss(), operand_dtype=torch.float32)
    xr = x.clone().requires_grad_(True)
    loss, d = fused(xr, tgt)
    loss.backward()
    np.testing.assert_allclose([d[k] for k in KEYS], gold[f"{name}/parts"], rtol=RTOL_FP32)
    np.testing.assert_allclose(lin.bias.grad.cpu().numpy(), gold[f"{name}/db"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(xr.grad[:8, :256].cpu().numpy(), gold[f"{name}/dx_head"], rtol=1e-3, atol=1e-7)
    np.testing.assert_allclose(lin.weight.grad[:64, :128].cpu().numpy(), gold[f"{name}/dw_head"], rtol=1e-3, atol=1e-7)
    with pytest.raises(_lib.YoloHeadError):
        linear_loss(x, w, b, tgt, grad_dtype=torch.bfloat16)      # fp32 operands take an fp32 gradient
    with pytest.raises(ValueError):
        linear_loss(x, w, b, tgt, precision="fp32")


def test_non_finite_activations_stay_in_their_row():
    """A NaN / inf in one image's activations makes that image's predictions (hence the total) non-finite, as torch's Linear
    + the reference's criterion would; every other image's gradient is what it was."""
    from yolo_v1_b200 import synth
    from yolo_v1_b200.head_linear import linear_loss

    M = 260
    g = torch.Generator(device="cuda").manual_seed(5)
    x = (torch.randn(M, 4096, device="cuda", generator=g) * 0.1).to(torch.bfloat16)
    w = (torch.randn(1470, 4096, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
    b = torch.rand(1470, device="cuda", generator=g) * 0.6
    tgt = synth.make_targets(M, seed=5, p_obj=0.2, size_mix=True).cuda()
    _, clean, _ = linear_loss(x, w, b, tgt, grad_dtype=torch.float32)
    for bad, row in ((float("nan"), 3), (float("inf"), 200)):
        xb = x.clone()
        xb[row, 77] = bad
        parts, grad, pred = linear_loss(xb, w, b, tgt, grad_dtype=torch.float32, want_pred=True)
        assert not torch.isfinite(parts[0])
        assert not torch.isfinite(pred[row]).all()
        keep = torch.ones(M, dtype=torch.bool, device="cuda")
        keep[row] = False
        assert torch.isfinite(pred[keep]).all() and torch.equal(grad[keep], clean[keep])


def test_second_backward_and_changing_upstream_gradient():
    """The gradient buffer is rescaled in place by upstream / loss_scale; a second backward over a retained graph, with
    another upstream factor, must not compound the first."""
    from yolo_v1_b200 import FusedHeadLoss, YOLOLoss, synth

    M = 200
    g = torch.Generator(device="cuda").manual_seed(8)
    x = torch.randn(M, 4096, device="cuda", generator=g) * 0.1
    tgt = synth.make_targets(M, seed=8, p_obj=0.1, size_mix=True).cuda()
    lin = torch.nn.Linear(4096, 1470).cuda()
    with torch.no_grad():
        lin.bias.add_(0.4)
    fused = FusedHeadLoss(lin, YOLOLoss(), operand_dtype=torch.bfloat16, backward_dtype=torch.float32)
    xr = x.clone().requires_grad_(True)
    loss, _ = fused(xr, tgt)
    loss.backward(retain_graph=True)
    g1 = xr.grad.clone()
    xr.grad = None
    (loss * 3.0).backward(retain_graph=True)
    g3 = xr.grad.clone()
    xr.grad = None
    loss.backward()
    torch.testing.assert_close(g3, 3.0 * g1, rtol=1e-4, atol=1e-7)   # fp32 GEMMs of g and of 3 g: sums round differently
    torch.testing.assert_close(xr.grad, g1, rtol=1e-4, atol=1e-7)
