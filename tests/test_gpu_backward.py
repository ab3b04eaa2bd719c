"""GPU parity of the fused backward kernels (run on the B200 box), through loss.backward() on the module API:
(1) against gradients torch.autograd produced on the live reference modules (committed fixtures, FlowLoss),
(2) against the fp64 numpy oracle's hand-derived backward at mid sizes with random output gradients,
(3) block-chunked backward (the gradient all-reduce overlap path) == single-call backward.

Tolerances: max-abs error relative to the tensor's max-abs, per gradient tensor.
  precision="fp32" (split-bf16): <= 2e-4 against the exact reference (fixtures / fp64 oracle).
  precision="bf16"             : <= 2e-2 against the oracle run with the SAME bf16 operand rounding
      (oracle.gemm_rounding("bf16")).  Against the exact fp32 reference a bf16 forward is not a meaningful
      per-tensor yardstick for gradients: LeakyReLU' is discontinuous, units whose pre-activation lies within
      the bf16 GEMM error of zero flip their mask and move single-sample contributions by O(1) (the CPU
      emulation shows the same 0.2-0.6 per-tensor deviations at the tiny test sizes).  The stated bound of the
      bf16 variant vs the exact reference is therefore global: rel-L2 over ALL gradients <= 0.1 at these sizes.
"""
import numpy as np
import pytest
import torch

import ipoke_inn_b200 as P
from oracle import inn_oracle as O
from conftest import load_golden, max_rel
from test_host_logic import CFG, FULL_CFG, build

pytestmark = pytest.mark.gpu

GTOL = {"fp32": 2e-4, "bf16": 2e-2}
GLOBAL_L2_BF16 = 0.1


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _load(name, precision):
    sd_ref, d = load_golden(name)
    net = build(name, precision=precision)
    net.load_state_dict({k: torch.from_numpy(v) for k, v in sd_ref.items()}, strict=True)
    return net.to(_dev()), sd_ref, d


def _report(errs, tol):
    bad = sorted(((e, k) for k, e in errs.items() if not (e < tol)), reverse=True)
    msg = "\n".join(f"  {k}: {e:.3e}" for e, k in bad[:12])
    assert not bad, f"{len(bad)}/{len(errs)} gradient tensors above {tol:g}:\n{msg}"


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("name", ["cfgS_cond", "uncond_small"])
def test_backward_matches_reference_autograd_fixture(name, precision):
    net, sd_ref, d = _load(name, precision)
    dev = _dev()
    x = torch.from_numpy(d["x"]).to(dev).requires_grad_(True)
    cond = torch.from_numpy(d["cond"]).to(dev).requires_grad_(True) if "cond" in d else None
    z, ld = net(x, cond) if cond is not None else net(x)
    loss, log = P.FlowLoss()(z, ld)
    loss.backward()
    tol = GTOL[precision]
    assert abs(float(loss.detach()) - float(d["loss"])) < tol * max(1.0, abs(float(d["loss"])))
    if precision == "bf16":
        # global deviation from the exact reference gradients (see module docstring)
        num = sum(float(((p.grad.cpu().numpy() - d["grad::" + k]) ** 2).sum()) for k, p in net.named_parameters())
        den = sum(float((d["grad::" + k] ** 2).sum()) for k, p in net.named_parameters())
        assert np.sqrt(num / den) < GLOBAL_L2_BF16
        assert max_rel(x.grad.cpu().numpy(), d["dx"]) < 5e-2
        return
    errs = {"dx": max_rel(x.grad.cpu().numpy(), d["dx"])}
    if cond is not None:
        errs["dcond"] = max_rel(cond.grad.cpu().numpy(), d["dcond"])
    for k, p in net.named_parameters():
        assert p.grad is not None, k
        errs[k] = max_rel(p.grad.cpu().numpy(), d["grad::" + k])
    _report(errs, tol)


KINK_MARGIN = 3e-5     # fp32-split forward error is ~1e-6 of a layer's rms: 30x clearance


def _oracle_case(name, B, precision, seed=5):
    """B samples whose hidden pre-activations all stay clear of LeakyReLU's kink (oracle.kink_distance): the
    derivative is discontinuous there, and with ~1e6 unit-samples a few would otherwise land within the
    forward rounding error of 0 and flip their mask relative to the fp64 oracle."""
    net, sd_ref, d = _load(name, precision)
    kind, kw = CFG[name]
    C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
    g = torch.Generator().manual_seed(seed)
    n = B + B // 2 + 8
    x = torch.randn(n, C, generator=g)
    cond = torch.randn(n, E, 1, 1, generator=g) if kind == "cond" else None
    dist = O.kink_distance(sd_ref, x.numpy(), None if cond is None else cond.numpy().reshape(n, -1))
    keep = torch.from_numpy(np.nonzero(dist > KINK_MARGIN)[0][:B])
    assert len(keep) == B, f"only {len(keep)} of {n} candidate samples clear the kink margin"
    x = x[keep].contiguous()
    cond = cond[keep].contiguous() if cond is not None else None
    dz = torch.randn(B, C, generator=g)
    dld = torch.randn(B, generator=g)
    return net, sd_ref, x, cond, dz, dld


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("name,B", [("cfgA_mini", 200), ("cfgB_mini", 130), ("cfgS_cond", 700), ("uncond_small", 1),
                                    ("cfgA_mini", 640), ("cfgB_mini", 1152)])
def test_backward_matches_fp64_oracle(name, B, precision):
    net, sd_ref, x, cond, dz, dld = _oracle_case(name, B, precision)
    _check_against_oracle(net, sd_ref, x, cond, dz, dld, precision)


# ~2e5 (cfg A) / ~1e5 (cfg B) hidden units per sample: a unit sits within m * rms of the kink with probability 0.8 m, so
# only exp(-0.8 m units) of the samples are clear: 4 % (cfg A) / 21 % (cfg B) at this margin -- candidates are drawn until
# enough are found (selection in fp32, row-chunked; the verification below is the fp64 oracle)
FULL_KINK_MARGIN = 2e-5


def _full_case(name, B, precision, seed=5, margin=FULL_KINK_MARGIN):
    """A full benchmark shape (no fixture): weights from the seed, ActNorm data-initialised on the device, then B
    kink-clear samples picked with the oracle on the resulting state_dict."""
    dev = _dev()
    net = build(name, precision=precision).to(dev)
    kind, kw = FULL_CFG[name]
    C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
    g = torch.Generator().manual_seed(seed)
    x0 = torch.randn(256, C, generator=g)
    c0 = torch.randn(256, E, 1, 1, generator=g) if kind == "cond" else None
    with torch.no_grad():
        net(x0.to(dev), c0.to(dev)) if c0 is not None else net(x0.to(dev))      # ActNorm data init
    sd_ref = O.state_dict_to_numpy(net.state_dict())
    xs, cs, found = [], [], 0
    for _ in range(40):
        x = torch.randn(512, C, generator=g)
        cond = torch.randn(512, E, 1, 1, generator=g) if kind == "cond" else None
        dist = O.kink_distance(sd_ref, x.numpy(), None if cond is None else cond.numpy().reshape(512, -1),
                               dtype=np.float32)
        keep = torch.from_numpy(np.nonzero(dist > margin)[0])
        xs.append(x[keep])
        if cond is not None:
            cs.append(cond[keep])
        found += len(keep)
        if found >= B:
            break
    assert found >= B, f"only {found} candidate samples clear the kink margin"
    x = torch.cat(xs)[:B].contiguous()
    cond = torch.cat(cs)[:B].contiguous() if kind == "cond" else None
    return net, sd_ref, x, cond, torch.randn(B, C, generator=g), torch.randn(B, generator=g)


@pytest.mark.parametrize("name,B", [("cfgA", 64), ("cfgB", 256)])
def test_backward_matches_fp64_oracle_at_the_benchmark_shapes(name, B):
    """The shapes whose training-step timings are quoted (H = 1536 / 1024: K-wide dgrad, 148-CTA wgrad, EPI_DU)."""
    net, sd_ref, x, cond, dz, dld = _full_case(name, B, "fp32")
    _check_against_oracle(net, sd_ref, x, cond, dz, dld, "fp32", stray_flips=True)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_large_batch_backward_rows_and_linearity(precision):
    """cfg A at 4096 rows (multi-wave dgrad / wgrad with K = batch = 4096): dx / dcond of 48 kink-clear rows embedded
    in the batch equal the fp64 oracle's (rows are independent), and the parameter gradients equal the sum of the
    gradients of the 8 row slices of 512 computed separately (linearity in the batch)."""
    B, nsub = 4096, 48
    net, sd_ref, xs, cs, dzs, dlds = _full_case("cfgA", nsub, precision)
    dev = _dev()
    g = torch.Generator().manual_seed(9)
    x = torch.randn(B, 128, generator=g)
    cond = torch.randn(B, 128, 1, 1, generator=g)
    dz = torch.randn(B, 128, generator=g)
    dld = torch.randn(B, generator=g)
    rows = torch.arange(nsub) * 85 + 3
    x[rows], cond[rows], dz[rows], dld[rows] = xs, cs, dzs, dlds

    def grads_of(lo, hi):
        net.zero_grad(set_to_none=True)
        xg = x[lo:hi].to(dev).requires_grad_(True)
        cg = cond[lo:hi].to(dev).requires_grad_(True)
        z, ld = net(xg, cg)
        ((z.reshape(hi - lo, -1) * dz[lo:hi].to(dev)).sum() + (ld * dld[lo:hi].to(dev)).sum()).backward()
        return xg.grad.cpu(), cg.grad.cpu().reshape(hi - lo, -1), [p.grad.detach().clone() for p in net.parameters()]

    dx, dc, full = grads_of(0, B)
    if precision == "fp32":
        with O.gemm_rounding(None):
            _, dx_ref, dc_ref = O.flow_backward(sd_ref, xs.numpy(), cs.numpy().reshape(nsub, -1), dzs.numpy(), dlds.numpy())
        assert max_rel(dx[rows].numpy(), dx_ref) < GTOL["fp32"]
        assert max_rel(dc[rows].numpy(), dc_ref) < GTOL["fp32"]
        from conftest import rel_l2
        assert rel_l2(dx[rows].numpy(), dx_ref) < 1e-4 and rel_l2(dc[rows].numpy(), dc_ref) < 1e-4
    acc = None
    for lo in range(0, B, 512):
        _, _, gs = grads_of(lo, lo + 512)
        acc = gs if acc is None else [a + b for a, b in zip(acc, gs)]
    # the slices run other tile shapes than the whole batch, and these 4096 rows are not kink-filtered: a unit whose
    # pre-activation lies within the forward rounding error of 0 flips its LeakyReLU mask between the two runs and moves
    # single-sample contributions (per-tensor maxima up to 1e-2 at ~2e5 units x 4096 samples); the batch sum as a whole
    # must agree -- a tiling or indexing bug in the multi-wave dgrad / wgrad would not stay below 1e-2 globally
    num = sum(float(((a - b).double() ** 2).sum()) for a, b in zip(full, acc))
    den = sum(float((b.double() ** 2).sum()) for b in acc)
    print(f"large-batch linearity, {precision}: global rel-L2 {np.sqrt(num / den):.3e}")
    assert np.sqrt(num / den) < (2e-3 if precision == "fp32" else 2e-2)


def _check_against_oracle(net, sd_ref, x, cond, dz, dld, precision, stray_flips=False):
    B = x.shape[0]
    dev = _dev()
    xg = x.to(dev).requires_grad_(True)
    cg = cond.to(dev).requires_grad_(True) if cond is not None else None
    z, ld = net(xg, cg) if cg is not None else net(xg)
    obj = (z.reshape(B, -1) * dz.to(dev)).sum() + (ld * dld.to(dev)).sum()
    obj.backward()
    with O.gemm_rounding("bf16" if precision == "bf16" else None):
        grads, dx_ref, dcond_ref = O.flow_backward(sd_ref, x.numpy(),
                                                   None if cond is None else cond.numpy().reshape(B, -1),
                                                   dz.numpy(), dld.numpy())
    tol = GTOL[precision]
    errs = {"dx": max_rel(xg.grad.cpu().numpy(), dx_ref)}
    if cg is not None:
        errs["dcond"] = max_rel(cg.grad.cpu().numpy().reshape(B, -1), dcond_ref)
    for k, p in net.named_parameters():
        errs[k] = max_rel(p.grad.cpu().numpy(), np.asarray(grads[k]).reshape(p.shape))
    if precision == "bf16":
        # two bf16 arithmetics differ by accumulation order: an activation that lands on a bf16 rounding boundary
        # moves by one ulp and can carry a unit across the kink; a handful of such chains is expected, a tiling
        # or indexing bug would break most tensors
        num = sum(float(((p.grad.cpu().numpy().astype(np.float64) - np.asarray(grads[k]).reshape(p.shape)) ** 2).sum())
                  for k, p in net.named_parameters())
        den = sum(float((np.asarray(grads[k]) ** 2).sum()) for k, p in net.named_parameters())
        assert np.sqrt(num / den) < 2e-2, f"global rel-L2 {np.sqrt(num / den):.3e}"
        n_bad = sum(1 for e in errs.values() if not e < tol)
        assert n_bad <= max(2, len(errs) // 10), f"{n_bad}/{len(errs)} tensors above {tol}"
        return
    if stray_flips:
        # reduced kink clearance (full shapes): a stray unit may still flip its LeakyReLU mask in one sample, which moves
        # a few entries of the tensors behind it; a tiling / indexing bug would break most tensors and the global norm
        num = sum(float(((p.grad.cpu().numpy().astype(np.float64) - np.asarray(grads[k]).reshape(p.shape)) ** 2).sum())
                  for k, p in net.named_parameters())
        den = sum(float((np.asarray(grads[k]) ** 2).sum()) for k, p in net.named_parameters())
        assert np.sqrt(num / den) < 2e-4, f"global rel-L2 {np.sqrt(num / den):.3e}"
        bad = {k: e for k, e in errs.items() if not e < tol}
        assert len(bad) <= max(2, len(errs) // 50), bad
        return
    _report(errs, tol)


def test_chunked_backward_equals_single_call_and_accumulates():
    name, B = "cfgS_cond", 96
    net, sd_ref, x, cond, dz, dld = _oracle_case(name, B, "fp32")
    dev = _dev()

    def run(hook):
        net.zero_grad(set_to_none=True)
        net.flow.engine.grad_ready_hook = hook
        xg = x.to(dev).requires_grad_(True)
        cg = cond.to(dev).requires_grad_(True)
        z, ld = net(xg, cg)
        ((z.reshape(B, -1) * dz.to(dev)).sum() + (ld * dld.to(dev)).sum()).backward()
        return ([p.grad.clone() for p in net.parameters()], xg.grad.clone(), cg.grad.clone())

    calls = []
    g0, dx0, dc0 = run(None)
    g1, dx1, dc1 = run(lambda eng, flat, kb, ke: calls.append((kb, ke)))
    net.flow.engine.grad_ready_hook = None
    assert calls == [(6, 8), (4, 6), (2, 4), (0, 2)]
    assert torch.allclose(dx0, dx1, rtol=0, atol=1e-6 * float(dx0.abs().max()))
    assert torch.allclose(dc0, dc1, rtol=0, atol=1e-6 * float(dc0.abs().max()))
    for a, b in zip(g0, g1):
        assert torch.allclose(a, b, rtol=0, atol=1e-6 * float(a.abs().max()) + 1e-12)
    # a second backward without zero_grad accumulates (autograd semantics)
    xg = x.to(dev).requires_grad_(True)
    z, ld = net(xg, cond.to(dev))
    ((z.reshape(B, -1) * dz.to(dev)).sum() + (ld * dld.to(dev)).sum()).backward()
    for a, p in zip(g1, net.parameters()):
        assert torch.allclose(p.grad, 2 * a, rtol=0, atol=2e-6 * float(a.abs().max()) + 1e-12)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_pair_wgrad_and_dgrad_equal_the_single_cta_kernels(precision):
    """The CTA-pair weight-gradient (256 x 256 tiles) and dgrad kernels against the single-CTA kernels the oracle tests
    above cover: same products, fp32 accumulation over the batch in both."""
    import os
    cfg = dict(flow_in_channels=64, flow_mid_channels_factor=16, flow_hidden_depth=1, n_flows=2,
               flow_conditioning_option="none", flow_embedding_channels=64)      # H = 1024; the 'cond' block runs G = 4 nets
    dev, B = _dev(), 2048
    g = torch.Generator().manual_seed(11)
    x = torch.randn(B, 64, generator=g).to(dev)
    cond = torch.randn(B, 64, 1, 1, generator=g).to(dev)
    dz, dld = torch.randn(B, 64, generator=g).to(dev), torch.randn(B, generator=g).to(dev)
    grads = {}
    for pair in ("1", "0"):
        os.environ["IPOKE_INN_PAIR_WGRAD"] = pair                 # read at plan creation (first call)
        os.environ["IPOKE_INN_PAIR_DGRAD"] = pair
        os.environ["IPOKE_INN_PAIR"] = "2"                        # both runs: pairs for every multi-wave launch (same forward)
        try:
            torch.manual_seed(42)
            net = P.SupervisedTransformer(dict(cfg), precision=precision).to(dev)
            with torch.no_grad():
                net(x, cond)                                       # ActNorm data init (identical for both)
            xg = x.clone().requires_grad_(True)
            z, ld = net(xg, cond)
            ((z.reshape(B, -1) * dz).sum() + (ld * dld).sum()).backward()
            grads[pair] = {k: p.grad.detach().cpu().numpy().copy() for k, p in net.named_parameters()}
            grads[pair]["dx"] = xg.grad.cpu().numpy().copy()
        finally:
            for k in ("IPOKE_INN_PAIR_WGRAD", "IPOKE_INN_PAIR_DGRAD", "IPOKE_INN_PAIR"):
                del os.environ[k]
    num = sum(float(((grads["1"][k].astype(np.float64) - grads["0"][k]) ** 2).sum()) for k in grads["1"])
    den = sum(float((grads["0"][k].astype(np.float64) ** 2).sum()) for k in grads["1"])
    # the forward passes are identical; the backward kernels differ in tile shape only (fp32 accumulation order, and in
    # bf16 a last-bit difference of a dgrad output can round differently downstream)
    assert np.sqrt(num / den) < (1e-5 if precision == "fp32" else 5e-3), np.sqrt(num / den)
    for k in grads["1"]:
        assert np.isfinite(grads["1"][k]).all(), k
        assert max_rel(grads["1"][k], grads["0"][k]) < (2e-5 if precision == "fp32" else 2e-2), k


def test_backward_graph_replay_equals_plain_launches():
    """A steady training loop hands the backward the same pointers every step (caching allocator): from the second
    step on it replays a captured CUDA graph.  Gradients of replayed steps must equal those of plain launches
    (IPOKE_INN_BWD_GRAPH=0) -- same kernels, same order; fp32 atomics (log-det, split-batch wgrad) allow last-bit noise."""
    import os
    name, B = "cfgA_mini", 96
    dev = _dev()
    g = torch.Generator().manual_seed(21)
    kind, kw = CFG[name]
    x = torch.randn(B, kw["flow_in_channels"], generator=g).to(dev)
    cond = torch.randn(B, kw["flow_embedding_channels"], 1, 1, generator=g).to(dev)
    dz, dld = torch.randn(B, kw["flow_in_channels"], generator=g).to(dev), torch.randn(B, generator=g).to(dev)
    res = {}
    for mode in ("1", "0"):
        os.environ["IPOKE_INN_BWD_GRAPH"] = mode
        try:
            net, _, _ = _load(name, "fp32")
            steps = []
            for it in range(4):
                net.zero_grad(set_to_none=True)
                xg = x.clone().requires_grad_(True)
                z, ld = net(xg, cond)
                ((z.reshape(B, -1) * dz).sum() + (ld * dld).sum()).backward()
                steps.append({k: p.grad.detach().cpu().numpy().copy() for k, p in net.named_parameters()})
                steps[-1]["__dx"] = xg.grad.detach().cpu().numpy().copy()
            res[mode] = steps
        finally:
            del os.environ["IPOKE_INN_BWD_GRAPH"]
    for it in range(4):
        for k in res["1"][it]:
            assert max_rel(res["1"][it][k], res["0"][it][k]) < 1e-5, (it, k)
            assert max_rel(res["1"][it][k], res["1"][0][k]) < 1e-5, (it, k)


def test_direct_param_grads_equal_autograd_accumulation():
    """engine.direct_param_grads = True assigns the flat-buffer views to p.grad inside backward(): same gradients as
    the AccumulateGrad route after one backward, and the same sums after a second backward without zero_grad."""
    name, B = "cfgS_cond", 40
    dev = _dev()
    g = torch.Generator().manual_seed(5)
    kind, kw = CFG[name]
    x = torch.randn(B, kw["flow_in_channels"], generator=g).to(dev)
    cond = torch.randn(B, kw["flow_embedding_channels"], 1, 1, generator=g).to(dev)
    res = {}
    for direct in (False, True):
        net, _, _ = _load(name, "fp32")
        net.flow.engine.direct_param_grads = direct
        out = []
        for it in range(2):                                   # second pass accumulates into the existing .grad
            xg = x.clone().requires_grad_(True)
            z, ld = net(xg, cond)
            loss, _ = P.FlowLoss()(z, ld)
            loss.backward()
            out.append({k: p.grad.detach().cpu().numpy().copy() for k, p in net.named_parameters()})
            out[-1]["__dx"] = xg.grad.detach().cpu().numpy().copy()
        res[direct] = out
    for it in range(2):
        for k in res[False][it]:
            assert max_rel(res[True][it][k], res[False][it][k]) < 1e-5, (it, k)
    k0 = next(k for k in res[True][0] if k != "__dx")
    assert max_rel(res[True][1][k0], 2 * res[True][0][k0]) < 1e-4      # really accumulated


def test_train_lists_carry_gradients_and_graph_can_be_retained():
    """the wrappers' train=True return (reference INN.py:80-81) must reach the parameters through its last
    entries, and retain_graph=True / a second backward must work."""
    net, sd_ref, d = _load("cfgS_cond", "fp32")
    dev = _dev()
    x = torch.from_numpy(d["x"]).to(dev)
    cond = torch.from_numpy(d["cond"]).to(dev)
    outs, lds = net(x, cond, train=True)
    loss = 0.5 * (outs[-1].reshape(x.shape[0], -1) ** 2).sum(1).mean() - lds[-1].mean()
    assert loss.requires_grad and not outs[0].requires_grad
    loss.backward(retain_graph=True)
    g1 = [p.grad.clone() for p in net.parameters()]
    net.zero_grad(set_to_none=True)
    loss.backward()
    for a, p in zip(g1, net.parameters()):
        assert torch.allclose(a, p.grad, rtol=1e-5, atol=1e-7 * float(a.abs().max()) + 1e-12)
    z, ld = net(x, cond)
    ref, _ = P.FlowLoss()(z, ld)
    assert abs(float(ref.detach()) - float(loss.detach())) < 1e-5 * max(1.0, abs(float(loss.detach())))
