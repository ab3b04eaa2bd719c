"""CPU tests of the host side: C-ABI surface, state_dict / RNG parity with the reference fixtures, API shape.
No kernel is launched here (no GPU in the build container)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

import ipoke_inn_b200 as P
from ipoke_inn_b200 import _capi
from conftest import ROOT, load_golden

CFG = {
    "cfgS_cond": ("cond", dict(flow_in_channels=8, flow_mid_channels_factor=8, flow_hidden_depth=2, n_flows=8,
                               flow_conditioning_option="none", flow_embedding_channels=16)),
    "uncond_small": ("uncond", dict(flow_in_channels=16, flow_mid_channels=64, flow_hidden_depth=2, n_flows=4)),
    "cfgA_mini": ("cond", dict(flow_in_channels=128, flow_mid_channels_factor=1, flow_hidden_depth=3, n_flows=2,
                               flow_conditioning_option="none", flow_embedding_channels=128)),
    "cfgB_mini": ("uncond", dict(flow_in_channels=512, flow_mid_channels=64, flow_hidden_depth=2, n_flows=2)),
}
# the full benchmark shapes: no fixtures, weights from the seed
FULL_CFG = {
    "cfgA": ("cond", dict(flow_in_channels=128, flow_mid_channels_factor=12, flow_hidden_depth=3, n_flows=8,
                          flow_conditioning_option="none", flow_embedding_channels=128)),
    "cfgB": ("uncond", dict(flow_in_channels=512, flow_mid_channels=1024, flow_hidden_depth=2, n_flows=8)),
}


def build(name, seed=42, **extra):
    kind, kw = (CFG.get(name) or FULL_CFG[name])
    torch.manual_seed(seed)
    if kind == "cond":
        return P.SupervisedTransformer(dict(kw), **extra)
    return P.UnsupervisedTransformer3(**kw, **extra)


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "ipoke_inn.h")).read()
    declared = set(re.findall(r"\b(ipoke_inn_[a-z_0-9]+)\s*\(", header))
    declared -= {"ipoke_inn_desc", "ipoke_inn_plan"}
    assert declared, "no declarations parsed"
    L = ctypes.CDLL(_capi.LIB_PATH)
    for sym in sorted(declared):
        assert hasattr(L, sym), f"{sym} declared in include/ipoke_inn.h but not exported"
    assert set(_capi.SYMBOLS) == declared
    assert _capi.lib().ipoke_inn_version() == 111


def test_plan_layout_without_gpu():
    L = _capi.lib()
    d = _capi.InnDesc(C=128, E=128, H=1536, depth=3, n_flows=8, conditional=1, precision=1, alpha=0.9, mlp_slope=0.01)
    plan = ctypes.c_void_p()
    assert L.ipoke_inn_plan_create(ctypes.byref(d), 0, ctypes.byref(plan)) == 0
    assert L.ipoke_inn_num_linears(plan) == 8 * 4 * 5 == 160          # 160 GEMMs / direction
    pb = L.ipoke_inn_packed_bytes(plan)
    assert 475e6 < pb < 500e6                                          # 237.5 M weights in bf16 + padding
    ws1, ws2 = L.ipoke_inn_workspace_bytes(plan, 64, 0), L.ipoke_inn_workspace_bytes(plan, 128, 0)
    assert 0 < ws1 < ws2
    out = (ctypes.c_longlong * 20)()
    assert L.ipoke_inn_debug_layout(plan, 64, out, 20) == 0
    assert out[10] == 64 + 128 and out[12] == 1536 and out[15] == 64  # ldU, Hp, Np
    L.ipoke_inn_plan_destroy(plan)
    d.precision = 0
    assert L.ipoke_inn_plan_create(ctypes.byref(d), 0, ctypes.byref(plan)) == 0
    assert 950e6 < L.ipoke_inn_packed_bytes(plan) < 1000e6             # hi + lo
    L.ipoke_inn_plan_destroy(plan)


@pytest.mark.parametrize("bad", [dict(C=7), dict(C=0), dict(H=0), dict(n_flows=0), dict(E=0), dict(precision=5)])
def test_plan_create_rejects_bad_shapes(bad):
    L = _capi.lib()
    kw = dict(C=8, E=16, H=64, depth=2, n_flows=8, conditional=1, precision=0, alpha=0.9, mlp_slope=0.01)
    kw.update(bad)
    plan = ctypes.c_void_p()
    rc = L.ipoke_inn_plan_create(ctypes.byref(_capi.InnDesc(**kw)), 0, ctypes.byref(plan))
    assert rc in (-1, -2)
    assert b"ipoke_inn" in L.ipoke_inn_last_error_string(rc)
    assert L.ipoke_inn_plan_create(None, 0, ctypes.byref(plan)) == -1


def test_null_arguments_are_errors_not_crashes():
    L = _capi.lib()
    assert L.ipoke_inn_forward(None, None, None, None, None, None, None, None, 4, 0, 1, None, None) == -1
    assert L.ipoke_inn_reverse(None, None, None, None, None, 4, 0, 1, None, None) == -1
    assert L.ipoke_inn_actnorm_stats(None, 4, 8, None, None) == -1
    assert L.ipoke_inn_flow_loss(None, None, 4, 8, 1.0, 1.0, None, None, None, None) == -1
    assert L.ipoke_inn_packed_bytes(None) == 0


@pytest.mark.parametrize("name", list(CFG))
def test_state_dict_and_rng_parity_with_reference(name):
    """Same seed => identical keys, shapes, dtypes, weights and (bit-exact) permutations as the reference."""
    sd_ref, _ = load_golden(name)
    net = build(name)
    sd = net.state_dict()
    assert list(sd.keys()) == list(sd_ref.keys())
    for k, v in sd.items():
        r = sd_ref[k]
        assert tuple(v.shape) == tuple(r.shape), k
        assert str(v.dtype).replace("torch.", "") == str(r.dtype), k
        if "norm_layer" in k:
            continue                       # fixture is post data-init; fresh modules have loc=0, scale=1, flag 0
        assert np.array_equal(v.numpy(), r), k
    net.load_state_dict({k: torch.from_numpy(v) for k, v in sd_ref.items()}, strict=True)
    assert int(net.flow.sub_layers[0].norm_layer.initialized) == 1


def test_wrapper_api_surface():
    net = build("cfgS_cond")
    assert net.emb_channels == 16 and net.in_channels == 8 and net.config["n_flows"] == 8
    f = net.flow
    assert (f.in_channels, f.cond_channels, f.mid_channels, f.num_blocks, f.n_flows) == (8, 16, 64, 2, 8)
    assert f.conditioning_option == "none" and f.reshape == "none"
    blk = f.sub_layers[0]
    for attr in ("norm_layer", "coupling", "activation", "shuffle"):
        assert hasattr(blk, attr)
    assert type(f.sub_layers[0].coupling).__name__ == "ConditionalDoubleVectorCouplingBlock_cond"
    assert type(f.sub_layers[1].coupling).__name__ == "ConditionalDoubleVectorCouplingBlock"
    assert net.get_last_layer().shape == (4, 64)
    assert f.last_outs == [] and f.last_logdets == []
    u = build("uncond_small")
    assert u.get_last_layer().shape == (8, 64) and not hasattr(u.flow.sub_layers[0], "activation")
    # extra kwargs tolerated like the reference (**config['architecture'] carries unrelated keys)
    P.UnsupervisedTransformer3(flow_in_channels=4, flow_mid_channels=8, flow_hidden_depth=1, n_flows=1,
                               base_distribution="normal", foo=1)


def test_no_cpu_fallback():
    net = build("cfgS_cond")
    x, c = torch.randn(4, 8), torch.randn(4, 16, 1, 1)
    with torch.no_grad():
        with pytest.raises(RuntimeError, match="CUDA"):
            net(x, c)
        with pytest.raises(RuntimeError, match="CUDA"):
            net(x, c, reverse=True)
        with pytest.raises(RuntimeError, match="fused"):
            net.flow.sub_layers[0].norm_layer(x)
        with pytest.raises(AssertionError):
            net.flow(x, c.reshape(4, 16))                  # reference asserts 4-D conditioning
    with pytest.raises(NotImplementedError):
        P.ConditionalFlow(8, 16, 64, 2, 2, conditioning_option="parallel")
    with pytest.raises(NotImplementedError):
        P.UnsupervisedTransformer3(flow_in_channels=7, flow_mid_channels=8, flow_hidden_depth=1, n_flows=1)


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_capi, "_lib", None)
    monkeypatch.setattr(_capi, "LIB_PATH", "/nonexistent/libipoke_inn.so")
    with pytest.raises(RuntimeError, match="no CPU or eager fallback"):
        _capi.lib()


def test_install_as_reference_shadows_import_paths():
    import sys
    P.install_as_reference()
    from models.modules.INN.INN import SupervisedTransformer, UnsupervisedTransformer3
    from models.modules.INN.loss import FlowLoss
    assert SupervisedTransformer is P.SupervisedTransformer and UnsupervisedTransformer3 is P.UnsupervisedTransformer3
    assert FlowLoss is P.FlowLoss
    from models.modules.INN.INN import SupervisedMacowTransformer, UnsupervisedMaCowTransformer3
    assert UnsupervisedMaCowTransformer3 is P.UnsupervisedMaCowTransformer3
    assert SupervisedMacowTransformer is P.SupervisedMacowTransformer
    for k in [k for k in sys.modules if k.startswith("models")]:
        del sys.modules[k]


def test_engine_parameter_walk_is_cached_and_follows_replaced_parameters():
    """engine._tensors() walks the module tree once; load_state_dict keeps the Parameter objects (cache stays valid),
    a rebuilt ActNorm parameter changes the key and triggers a new walk (and a new gradient layout)."""
    net = build("cfgS_cond")
    eng = net.flow.engine
    t0 = eng._tensors()
    assert eng._tensors() is t0
    net.load_state_dict(net.state_dict())
    assert eng._tensors() is t0
    lay0 = eng.grad_layout()
    blk = net.flow.sub_layers[3]
    blk.norm_layer.loc = torch.nn.Parameter(blk.norm_layer.loc.detach().clone())
    t1 = eng._tensors()
    assert t1 is not t0 and t1[2][3] is blk.norm_layer.loc
    assert eng.grad_layout() is not lay0
    assert set(id(p) for p in eng.grad_layout()[3]) == set(id(p) for p in net.flow.parameters())


def test_direct_param_grads_host_logic_with_a_stub_engine():
    """FlowFunction.backward with engine.direct_param_grads: the views of the flat buffer become p.grad on the first
    backward and are accumulated on the second; autograd itself receives None for the parameters.  The kernels are
    replaced by a stub engine that fills the flat buffer with a known pattern (CPU, host logic only)."""
    from ipoke_inn_b200.backward import FlowFunction

    params = [torch.nn.Parameter(torch.zeros(3, 2)), torch.nn.Parameter(torch.zeros(5))]

    class StubEngine:
        grad_ready_hook = None
        direct_param_grads = True
        last_flat_grad = None
        n_flows = 1

        def forward_train(self, x, cond):
            B = x.shape[0]
            z, ld = x * 2.0, torch.zeros(B)
            return z, ld, torch.zeros(1, B, x.shape[1]), torch.zeros(1, B), torch.zeros(1, dtype=torch.uint8), x, None

        def grad_layout(self):
            return None, 11, [(0, 11)], params

        def new_grad_buffer(self):
            flat = torch.arange(11, dtype=torch.float32)
            return flat, list(torch._utils._unflatten_dense_tensors(flat, params)), (None, None, None, None)

        def backward(self, x2, c2, outs, saved, dz, dld, ptrs, want_dx=True, want_dcond=True, **kw):
            return dz * 2.0, None

    eng = StubEngine()
    x = torch.randn(4, 6, requires_grad=True)
    for it in (1, 2):
        z, ld, _, _ = FlowFunction.apply(eng, x, None, *params)
        (z.sum() + ld.sum()).backward()
        assert torch.equal(params[0].grad, it * torch.arange(6, dtype=torch.float32).view(3, 2))
        assert torch.equal(params[1].grad, it * torch.arange(6, 11, dtype=torch.float32))
    assert torch.allclose(x.grad, torch.full((4, 6), 4.0))           # 2 backward passes x dz * 2
    assert eng.last_flat_grad is not None


def test_nvtx_ranges_are_opt_in():
    """IPOKE_NVTX=1 wraps the host entry points in torch.cuda.nvtx ranges; by default they are the plain functions."""
    import subprocess
    import sys
    from ipoke_inn_b200 import engine
    assert not hasattr(engine.FlowEngine.forward, "__wrapped__")
    code = ("import ipoke_inn_b200 as P; from ipoke_inn_b200 import engine, macow_flow, optim; "
            "assert hasattr(engine.FlowEngine.forward, '__wrapped__') and hasattr(engine.FlowEngine.backward, '__wrapped__'); "
            "assert hasattr(macow_flow.Program.run, '__wrapped__') and hasattr(optim.FusedAdam.step, '__wrapped__'); print('ok')")
    env = dict(os.environ, IPOKE_NVTX="1")
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env,
                         cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert res.returncode == 0 and res.stdout.strip().endswith("ok"), res.stderr[-500:]
