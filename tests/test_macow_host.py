"""CPU checks of the MaCow drop-in module (ipoke-main-opticalflow_b200/macow.py): state_dict keys and shapes, RNG draw
order, the exported C ABI (include/ipoke_macow.h), argument errors.  No compute without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

import ipoke_inn_b200 as P
from ipoke_inn_b200 import _capi, macow

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def golden(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    C, k0, k1, B, H, W = (int(v) for v in g["meta"])
    return g, dict(C=C, k=[k0, k1], B=B, H=H, W=W, act=str(g["activation"]), alpha=float(g["alpha"]))


@pytest.mark.parametrize("name", ["macow_elu", "macow_relu", "macow_lrelu"])
def test_state_dict_keys_shapes_and_rng_order(name):
    g, m = golden(name)
    torch.manual_seed(42)                                       # the seed tests/golden/make_macow_golden.py used
    unit = macow.MaCowUnit(m["C"], m["k"], activation=m["act"], alpha=m["alpha"])
    ref_keys = [k[4:] for k in g.files if k.startswith("sd::")]
    sd = unit.state_dict()
    assert list(sd.keys()) == ref_keys
    for k in ref_keys:
        assert tuple(sd[k].shape) == g["sd::" + k].shape, k
        assert sd[k].numpy().dtype == g["sd::" + k].dtype, k
        # tensors the reference's first forward / the fixture generator did not overwrite: identical draws
        if k.endswith("shift_conv.weight") or k.endswith("weight_v"):
            assert np.array_equal(sd[k].numpy(), g["sd::" + k]), k
    # a fresh weight-norm gain is ||v|| (nn.utils.weight_norm), so the fresh 1x1 weight equals v
    v = sd["conv1.net.conv1x1.conv.weight_v"]
    np.testing.assert_allclose(sd["conv1.net.conv1x1.conv.weight_g"].flatten(), v.flatten(1).norm(dim=1), rtol=1e-6)
    unit.load_state_dict({k: torch.from_numpy(g["sd::" + k]) for k in ref_keys})      # strict


def test_header_symbols_are_exported():
    hdr = open(os.path.join(ROOT, "include", "ipoke_macow.h")).read()
    names = sorted(set(re.findall(r"\b(ipoke_macow_[a-z_]+)\s*\(", hdr)))
    assert names == ["ipoke_macow_chanmap", "ipoke_macow_cond_bias", "ipoke_macow_cond_floats", "ipoke_macow_forward",
                     "ipoke_macow_launch_info", "ipoke_macow_nice", "ipoke_macow_nice_floats",
                     "ipoke_macow_nice_workspace_floats", "ipoke_macow_pack_nice", "ipoke_macow_pack_unit",
                     "ipoke_macow_pack_unit_cond", "ipoke_macow_reverse", "ipoke_macow_unit_floats",
                     "ipoke_macow_units_forward", "ipoke_macow_units_forward_cond", "ipoke_macow_units_reverse",
                     "ipoke_macow_units_reverse_cond"]
    L = _capi.lib()
    for n in names:
        assert hasattr(L, n), n


def test_unit_floats_and_launch_geometry_without_gpu():
    L = macow._lib()
    d = macow.MacowDesc(16, 64, 2, 3, 1, 1.0)
    per_flow = 16 * 2 * 3 * 64 + 64 * 32 + 32
    assert L.ipoke_macow_unit_floats(ctypes.byref(d)) == 4 * per_flow + 4 * 16
    smem, thr, tile = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    assert L.ipoke_macow_launch_info(ctypes.byref(d), 8, 8, 1, ctypes.byref(smem), ctypes.byref(thr), ctypes.byref(tile)) == 0
    # inverse: two image copies + one hidden line per channel group (512 threads / 32 work items, at most 4 groups)
    # ... + one flow's packed weights staged in shared memory
    assert tile.value == 1 and smem.value == (2 * 16 * 64 + 4 * 8 * 64 + per_flow) * 4 and thr.value % 32 == 0
    assert L.ipoke_macow_launch_info(ctypes.byref(d), 8, 8, 0, ctypes.byref(smem), ctypes.byref(thr), ctypes.byref(tile)) == 0
    assert tile.value == 8                                      # the whole 8 x 8 image's hidden tile fits the budget
    bad = macow.MacowDesc(16, 64, 2, 4, 1, 1.0)                 # even centred kernel: the reference asserts
    assert L.ipoke_macow_unit_floats(ctypes.byref(bad)) == 0
    huge = macow.MacowDesc(512, 512, 2, 3, 1, 1.0)              # 512 x 32 x 32 floats do not fit one SM
    assert L.ipoke_macow_launch_info(ctypes.byref(huge), 32, 32, 1, None, None, None) == -2


def test_unbuilt_variants_and_cpu_tensors_fail_loudly():
    with pytest.raises(NotImplementedError):
        macow.MaCowUnit(4, [2, 3], h_channels=8, cond_conv=True, cond_conv_hidden_channels=16)
    cond = macow.MaCowUnit(4, [2, 3], h_channels=8)                # conditional unit: 1x1 convolutions 16 + 8 wide
    assert cond.conv1.net.conv1x1.conv.weight_v.shape == (8, 24, 1, 1) and cond._desc.h_channels == 8
    L = macow._lib()
    assert L.ipoke_macow_cond_floats(ctypes.byref(cond._desc), 2) == 8 * 2 * 8 * 4          # h_channels x (units * 8 C)
    assert L.ipoke_macow_cond_floats(ctypes.byref(macow.MacowDesc(4, 16, 2, 3, 1, 1.0)), 2) == 0
    # conditional descriptor without its conditioning product (and the other way round) is refused before any launch
    assert L.ipoke_macow_units_forward(ctypes.byref(cond._desc), 1, 1, 1, 64, 1, 64, None, 0, 1, 4, 4, 0, 6, None) == -1
    with pytest.raises(NotImplementedError):
        macow.MaCowUnit(4, [2, 3], transform="nlsq")
    with pytest.raises(NotImplementedError):
        macow.MaCowUnit(4, [2, 3], p_dropout=0.1)
    with pytest.raises(AssertionError):
        macow.MaCowUnit(4, [2, 2])                               # even width: ShiftedConv2d's own assert
    unit = macow.MaCowUnit(4, [2, 3])
    with torch.no_grad(), pytest.raises(RuntimeError, match="CUDA"):
        unit(torch.randn(2, 4, 6, 6))
    assert "MaCowUnit" in P.__all__


def test_flow_sizes_argument_errors_and_unbuilt_variants_without_gpu():
    L = macow._lib()
    d = macow.NiceDesc(16, 0, 1, 16, 16, 1, 2048, 1, 1.0)        # continuous / up coupling of a 32-channel step
    assert L.ipoke_macow_nice_floats(ctypes.byref(d)) == 9 * 16 * 2048 + 2048 * 2048 + 9 * 2048 * 32 + 32
    # preferred workspace: as many samples as keep a pass (two hidden-activation buffers + tap outputs) within 48 MB
    per_sample = 64 * (2 * 2048 + 9 * 32)
    assert L.ipoke_macow_nice_workspace_floats(ctypes.byref(d), 4096, 8, 8) == (48 * 2 ** 20 // 4 // per_sample) * per_sample
    assert L.ipoke_macow_nice_workspace_floats(ctypes.byref(d), 3, 8, 8) == 3 * per_sample
    assert L.ipoke_macow_nice_floats(ctypes.byref(macow.NiceDesc(0, 0, 1, 16, 16, 1, 64, 1, 1.0))) == 0
    assert L.ipoke_macow_nice(ctypes.byref(d), None, None, 0, None, 0, 1, 8, 8, None, 0, None) == -1     # NULL pointers
    assert L.ipoke_macow_chanmap(None, 0, 1, 8, 8, None, 0, None, None, 0, 0, 0, 0, None, None) == -1
    cfg = dict(num_steps=[1], flow_in_channels=8, flow_mid_channels=16, factor=4, transform="affine",
               prior_transform="affine", kernel_size=[2, 3], coupling_type="conv", activation="elu")
    net = P.UnsupervisedMaCowTransformer3(cfg)
    assert net.flow.z_channels == 6 and net.flow._program is None
    assert [op[0] for op in net.flow.ops()] == ["glue", "units", "nice", "nice", "glue", "units", "nice", "nice",
                                                 "glue", "nice", "glue"]
    for bad in (dict(cfg, transform="nlsq"), dict(cfg, use1x1=True), dict(cfg, prior_transform="additive")):
        with pytest.raises(NotImplementedError):
            P.UnsupervisedMaCowTransformer3(bad)
    with pytest.raises(NotImplementedError):
        P.NICE2d(8, h_channels=4)
    ccfg = dict(cfg, h_channels=6, flow_attn_heads=4, cond_conv_hidden_channels=16)
    cnet = P.SupervisedMacowTransformer(ccfg)
    assert cnet.flow.h_channels == 6 and cnet.flow.layers[0][0].units1[0].h_channels == 6
    assert cnet.flow.layers[0][0].coupling1_up.net.conv1.weight.shape[1] == 4           # couplings: no conditioning
    for bad in (dict(ccfg, condition_nice=True), dict(ccfg, cond_conv=True)):
        with pytest.raises(NotImplementedError):
            P.SupervisedMacowTransformer(bad)
    with pytest.raises(KeyError):
        P.SupervisedMacowTransformer(dict(cfg, h_channels=6, flow_attn_heads=4))      # the reference requires the key too
