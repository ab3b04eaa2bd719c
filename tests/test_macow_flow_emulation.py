"""The full MaCow flow on the CPU: the Python host code of ipoke-main-opticalflow_b200/macow_flow.py driving a g++ build
of the DEVICE code (csrc/macow_kernel.cuh, csrc/macow_flow_kernel.cuh) and of the launch logic (csrc/macow_flow_host.h)
through tests/emulation/macow_flow_emul.cpp: every CUDA thread is a fiber, __syncthreads() a scheduler round, CTAs run
with the real thread counts / grids / shared-memory carving, in ascending and in descending thread order (a missing
barrier gives a wrong result in one of the two).  Checked against the fixtures of the live reference
(tests/golden/macow_stack_*.npz) and the float64 oracle.  The emulation library is handed to the host code by
monkeypatching; the package itself never loads it (CPU tensors raise in the product path -- asserted below)."""
import ctypes
import os
import subprocess

import numpy as np
import pytest
import torch

import ipoke_inn_b200 as P
from ipoke_inn_b200 import macow
from oracle import macow_stack_oracle as so
from conftest import max_rel

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.path.join(HERE, "emulation", "macow_flow_emul.cpp")
OUT = os.path.join(HERE, "emulation", "_build", "libmacow_flow_emul.so")
DEPS = [SRC, os.path.join(HERE, "emulation", "cuda_emul.h")] + [
    os.path.join(ROOT, "ipoke-main-opticalflow_b200", "csrc", f)
    for f in ("macow_kernel.cuh", "macow_flow_kernel.cuh", "macow_flow_host.h")] + [
    os.path.join(ROOT, "include", "ipoke_macow.h")]


@pytest.fixture(scope="module")
def emul_lib():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if not os.path.exists(OUT) or os.path.getmtime(OUT) < max(os.path.getmtime(f) for f in DEPS):
        subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-Wno-unknown-pragmas", "-Wno-attributes",
                        "-I", os.path.join(ROOT, "include"), "-o", OUT, SRC], check=True)
    return macow._bind(ctypes.CDLL(OUT))


@pytest.fixture()
def emulated(emul_lib, monkeypatch):
    """The host code's three touch points with CUDA, redirected: library, stream, the CUDA-only input check."""
    def check(x, channels):
        assert x.dim() == 4 and x.shape[1] == channels and x.dtype == torch.float32
        return x.contiguous()
    monkeypatch.setattr(macow, "_lib", lambda: emul_lib)
    monkeypatch.setattr(macow, "_stream", lambda dev: None)
    monkeypatch.setattr(macow, "_check_input", check)
    emul_lib.emul_set_descending(0)
    return emul_lib


def load(name):
    g = np.load(os.path.join(HERE, "golden", name + ".npz"))
    sd = {k[4:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("sd::")}
    C, mid, factor, k0, k1 = (int(v) for v in g["meta"])
    cfg = dict(num_steps=[int(v) for v in g["num_steps"]], flow_in_channels=C, flow_mid_channels=mid, factor=factor,
               transform="affine", prior_transform="affine", kernel_size=[k0, k1], coupling_type="conv",
               activation=str(g["activation"]))
    if "h_channels" in g.files and int(g["h_channels"]) > 0:      # SupervisedMacowTransformer reads these keys too
        cfg.update(h_channels=int(g["h_channels"]), flow_attn_heads=4, cond_conv_hidden_channels=16)
    return g, sd, cfg


def test_product_path_refuses_cpu_tensors():
    g, sd, cfg = load("macow_stack_a")
    net = P.UnsupervisedMaCowTransformer3(cfg)
    with torch.no_grad(), pytest.raises(RuntimeError, match="CUDA"):
        net(torch.from_numpy(g["x"]))


def test_construction_matches_the_reference_draw_for_draw():
    """Same seed -> the reference's constructor-time weights and permutations, same keys / shapes / dtypes."""
    for name in ("macow_stack_a", "macow_stack_b"):
        g, sd, cfg = load(name)
        torch.manual_seed(42)
        fresh = P.UnsupervisedMaCowTransformer3(cfg).state_dict()
        assert list(fresh.keys()) == list(sd.keys())
        for k, v in sd.items():
            assert fresh[k].shape == v.shape and fresh[k].dtype == v.dtype, k
            if k.endswith(("shuffle_idx", "weight_v", "conv1.weight", "conv2.weight", "shift_conv.weight")):
                assert torch.equal(fresh[k], v), k                     # untouched by the data-dependent initialisation


def test_first_call_initialisation(emulated):
    g, sd, cfg = load("macow_stack_b")
    torch.manual_seed(42)
    net = P.UnsupervisedMaCowTransformer3(cfg).eval()
    with torch.no_grad():
        out, ld = net(torch.from_numpy(g["x"]))
    assert max_rel(out.numpy(), g["init_out"]) <= 2e-5 and max_rel(ld.numpy(), g["init_logdet"]) <= 2e-5
    new = net.state_dict()
    for k in g.files:
        if k.startswith("init::"):
            np.testing.assert_allclose(new[k[6:]].numpy(), g[k], rtol=1e-4, atol=2e-5, err_msg=k)
    assert all(int(v) == 1 for k, v in new.items() if k.endswith("initialized"))
    with torch.no_grad():                                              # second call takes the plain program
        out2, ld2 = net(torch.from_numpy(g["x"]))
    assert max_rel(out2.numpy(), out.numpy()) <= 1e-6 and max_rel(ld2.numpy(), ld.numpy()) <= 1e-6


@pytest.mark.parametrize("name,descending", [("macow_stack_a", 0), ("macow_stack_b", 1)])
def test_flow_matches_reference_fixtures(emulated, name, descending):
    g, sd, cfg = load(name)
    net = P.UnsupervisedMaCowTransformer3(cfg).eval()
    net.load_state_dict(sd)
    emulated.emul_set_descending(descending)
    x = torch.from_numpy(g["x"])
    with torch.no_grad():
        out, ld = net(x)
        smp = net(torch.from_numpy(g["z_in"]), reverse=True)
    assert torch.equal(x, torch.from_numpy(g["x"]))                    # the caller's tensor is not modified
    assert max_rel(out.numpy(), g["out64"]) <= 2e-5 and max_rel(ld.numpy(), g["logdet64"]) <= 2e-5
    assert max_rel(out.numpy(), g["out"]) <= 2e-5
    assert max_rel(smp.numpy(), g["x_smp64"]) <= 2e-4 and max_rel(smp.numpy(), g["x_smp"]) <= 2e-4


def test_conditional_flow(emulated):
    """SupervisedMacowTransformer (INN.py:447-482): same-seed construction, first-call initialisation, forward / inverse
    with the conditioning, one conditional unit and one step on their own -- against the live-reference fixture."""
    g, sd, cfg = load("macow_stack_c")
    x, z, h = torch.from_numpy(g["x"]), torch.from_numpy(g["z_in"]), torch.from_numpy(g["h"])
    torch.manual_seed(42)
    net = P.SupervisedMacowTransformer(cfg).eval()
    fresh = net.state_dict()
    assert list(fresh.keys()) == list(sd.keys())
    for k, v in sd.items():
        assert fresh[k].shape == v.shape and fresh[k].dtype == v.dtype, k
        if k.endswith(("shuffle_idx", "weight_v", "conv1.weight", "conv2.weight", "shift_conv.weight")):
            assert torch.equal(fresh[k], v), k
    with torch.no_grad():
        out, ld = net(x, h)
    assert max_rel(out.numpy(), g["init_out"]) <= 2e-5 and max_rel(ld.numpy(), g["init_logdet"]) <= 2e-5
    for k in g.files:
        if k.startswith("init::"):
            np.testing.assert_allclose(net.state_dict()[k[6:]].numpy(), g[k], rtol=1e-4, atol=2e-5, err_msg=k)
    net.load_state_dict(sd)
    emulated.emul_set_descending(1)
    with torch.no_grad():
        out, ld = net(x, h)
        smp = net(z, h, reverse=True)
        assert max_rel(out.numpy(), g["out64"]) <= 2e-5 and max_rel(ld.numpy(), g["logdet64"]) <= 2e-5
        assert max_rel(smp.numpy(), g["x_smp64"]) <= 2e-4
        step = net.flow.layers[0][0]
        o, l = step.units1[0](x, h=h)
        assert max_rel(o.numpy(), g["unit_out"]) <= 2e-5 and max_rel(l.numpy(), g["unit_logdet"]) <= 2e-5
        assert max_rel(step.units1[0](z, h=h, reverse=True).numpy(), g["unit_rev"]) <= 2e-4
        o, l = step(x, h=h)
        assert max_rel(o.numpy(), g["step_out"]) <= 2e-5 and max_rel(l.numpy(), g["step_logdet"]) <= 2e-5
        assert max_rel(step(z, h=h, reverse=True).numpy(), g["step_rev"]) <= 2e-4
        # a different conditioning gives a different result; a missing or mis-shaped one is refused
        assert not torch.equal(net(x, h + 0.5)[0], out)
        with pytest.raises(ValueError):
            net.flow(x)
        with pytest.raises(ValueError):
            net(x, h[:, :3])


def test_components_called_like_the_reference_calls_them(emulated):
    """One NICE2d per split type / order, one step, one prior."""
    g, sd, cfg = load("macow_stack_a")
    net = P.UnsupervisedMaCowTransformer3(cfg).eval()
    net.load_state_dict(sd)
    emulated.emul_set_descending(1)
    x, z = torch.from_numpy(g["x"]), torch.from_numpy(g["z_in"])
    step = net.flow.layers[0][0]
    with torch.no_grad():
        for cn in ("coupling1_up", "coupling1_dn", "coupling2_up", "coupling2_dn"):
            o, l = getattr(step, cn)(x)
            assert max_rel(o.numpy(), g[cn + "_out"]) <= 2e-5 and max_rel(l.numpy(), g[cn + "_logdet"]) <= 2e-5, cn
            assert max_rel(getattr(step, cn)(z, reverse=True).numpy(), g[cn + "_rev"]) <= 2e-5, cn
        o, l = step(x)
        assert max_rel(o.numpy(), g["step_out"]) <= 2e-5 and max_rel(l.numpy(), g["step_logdet"]) <= 2e-5
        assert max_rel(step(z, reverse=True).numpy(), g["step_rev"]) <= 2e-4
        o, l = net.flow.priors[0](x)
        assert max_rel(o.numpy(), g["prior_out"]) <= 2e-5 and max_rel(l.numpy(), g["prior_logdet"]) <= 2e-5
        assert max_rel(net.flow.priors[0](z, reverse=True).numpy(), g["prior_rev"]) <= 2e-5


@pytest.mark.parametrize("Hd", [40, 100])
def test_coupling_rows_beyond_one_tile_and_workspace_passes(emulated, Hd):
    """A coupling on 3 x 7 x 9 pixels = 189 rows (two 128-row tiles, the second ragged) with hidden width 40 (not a
    multiple of the 16-deep K slice, one ragged 128-column tile) and a workspace of 2 samples (passes of 2 + 1); at
    hidden width 100 the tap product of the last convolution runs as two K slices (64 + 36) parked in the first
    activation buffer."""
    rng = np.random.default_rng(5)
    C, B, H, W = 6, 3, 7, 9
    nice = P.NICE2d(C, hidden_channels=Hd, split_type="continuous", order="down", factor=3, activation="leaky_relu",
                    alpha=0.8).eval()
    with torch.no_grad():
        nice.net.conv3.conv.weight_g.copy_(torch.from_numpy(0.2 + 0.3 * rng.random((4, 1, 1, 1)).astype(np.float32)))
        nice.net.conv3.conv.bias.copy_(torch.from_numpy(0.1 * rng.standard_normal(4).astype(np.float32)))
        nice.net.conv3.initialized.fill_(1)
    sd = {k: v.numpy() for k, v in nice.state_dict().items()}
    x = rng.standard_normal((B, C, H, W)).astype(np.float32)
    want, want_ld = so.nice_apply(sd, x, False, "continuous", 3, "leaky_relu", 0.8)
    d = nice._desc
    n = emulated.ipoke_macow_nice_floats(ctypes.byref(d))
    packed = np.zeros(n, dtype=np.float32)
    arrs = [p.detach().numpy() for p in nice.params()]
    ptrs = (ctypes.c_void_p * 5)(*[a.ctypes.data for a in arrs])
    assert emulated.ipoke_macow_pack_nice(ctypes.byref(d), ptrs, packed.ctypes.data, None) == 0
    per_sample = H * W * (2 * Hd + 9 * 4)                                # round4(2 c_t) = 4
    ws = np.zeros(2 * per_sample, dtype=np.float32)
    got, ld = x.copy(), np.full(B, 1.5, dtype=np.float32)
    assert emulated.ipoke_macow_nice(ctypes.byref(d), packed.ctypes.data, got.ctypes.data, C * H * W, ld.ctypes.data, 0,
                                     B, H, W, ws.ctypes.data, ws.size, None) == 0
    assert max_rel(got, want) <= 1e-5 and max_rel(ld - 1.5, want_ld) <= 1e-5
    back = got.copy()
    assert emulated.ipoke_macow_nice(ctypes.byref(d), packed.ctypes.data, back.ctypes.data, C * H * W, None, 1, B, H, W,
                                     ws.ctypes.data, ws.size, None) == 0
    assert max_rel(back, x) <= 1e-5
    # a workspace smaller than one sample is refused
    assert emulated.ipoke_macow_nice(ctypes.byref(d), packed.ctypes.data, back.ctypes.data, C * H * W, None, 1, B, H, W,
                                     ws.ctypes.data, per_sample - 1, None) != 0


def test_gpu_test_functions_run_through_the_emulation(emulated, monkeypatch):
    """The GPU parity tests themselves (tests/test_zy_gpu_macow_flow.py), device swapped for the CPU emulation: what the
    B200 run asserts is asserted here first (their larger shapes take minutes in the numpy oracle and stay GPU-only)."""
    import test_zy_gpu_macow_flow as G
    monkeypatch.setattr(G, "DEV", torch.device("cpu"))
    emulated.emul_set_descending(1)
    G.test_conditional_flow_reference_fixture()
    G.test_components_called_like_the_reference_calls_them()
    G.test_first_call_initialisation_matches_reference()
    G.test_against_float64_oracle(12, 20, 4, [2], 5, 7, 3, "leaky_relu", None, monkeypatch)


def test_copy_and_pickle_rebuild_the_program(emulated):
    """copy.deepcopy / torch.save(module) of a flow that has already run: the copy carries parameters and buffers, not the
    packed program, and computes the same result; broadcast-style writes through .data need invalidate()."""
    import copy
    import io
    g, sd, cfg = load("macow_stack_c")
    net = P.SupervisedMacowTransformer(cfg).eval()
    net.load_state_dict(sd)
    x, h = torch.from_numpy(g["x"]), torch.from_numpy(g["h"])
    with torch.no_grad():
        want, want_ld = net(x, h)
        assert net.flow._program is not None
        twin = copy.deepcopy(net)
        assert twin.flow._program is None and net.flow._program is not None
        buf = io.BytesIO()
        torch.save(net, buf)
        buf.seek(0)
        loaded = torch.load(buf, weights_only=False)
        for other in (twin, loaded):
            out, ld = other(x, h)
            assert torch.equal(out, want) and torch.equal(ld, want_ld)
        # a write through .data is invisible to the version keys until invalidate()
        unit = net.flow.layers[0][0].units1[0]
        unit.conv1.net.conv1x1.conv.bias.data.add_(0.5)
        stale, _ = net(x, h)
        assert torch.equal(stale, want)
        net.flow.invalidate()
        fresh, _ = net(x, h)
        assert not torch.equal(fresh, want)
