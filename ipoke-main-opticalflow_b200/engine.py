"""Host-side driver of one flow stack: owns the C-ABI plan, the packed parameter buffer and the scratch
workspaces, mirrors the ActNorm ``initialized`` flags on the host (no per-call ``.item()`` sync as in the
reference, modules.py:288) and exposes forward / reverse / backward on 2-D fp32 CUDA tensors."""
import ctypes

import torch

from . import _capi
from ._nvtx import annotate


def _ptr_array(tensors):
    arr = (ctypes.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


class FlowEngine:
    def __init__(self, flow, conditional, C, E, H, depth, n_flows, alpha, precision="fp32"):
        if precision not in _capi.PRECISIONS:
            raise ValueError(f"precision must be one of {list(_capi.PRECISIONS)}")
        if C % 2 != 0:
            raise NotImplementedError("ipoke_inn_b200 covers even flow_in_channels only (all reference configs)")
        self.flow = flow                      # not registered as a sub-module: FlowEngine is a plain object
        self.desc = _capi.InnDesc(C=C, E=E, H=H, depth=depth, n_flows=n_flows, conditional=int(conditional),
                                  precision=_capi.PRECISIONS[precision], alpha=float(alpha), mlp_slope=0.01)
        self.precision = precision
        self.C, self.E, self.n_flows = C, E, n_flows
        self._plan = None
        self._device = None
        self._packed = None
        self._packed_sig = None
        self._ws = {}
        self._init_epoch = None               # sum of ActNorm._state_epoch when `initialized` was last read
        self.assume_frozen = False            # skip the per-call parameter-version check (inference loops)
        self.sync_actnorm_init = True         # all-reduce ActNorm init moments when torch.distributed is up
        self.launches = 0
        self.bwd_launches = 0
        self.grad_ready_hook = None           # callable(engine, flat_grad, block_begin, block_end): see parallel.GradSync
        self.grad_chunk_blocks = 2
        self.last_flat_grad = None
        self.direct_param_grads = False       # True: backward() assigns / accumulates p.grad itself (see backward.py)
        self.flat_grads_only = False          # True: backward() leaves p.grad alone; the flat buffer (last_flat_grad) is the
                                              # only result (set by ShardedAdam, which consumes the buffer directly)
        self._grad_layout = None
        self.streams = 1
        self._pm = None                       # packed-master state (adopt_packed_master)

    # ------------------------------------------------------------------ plumbing
    def set_concurrency(self, streams):
        """Tell the plan that `streams` calls run concurrently on different CUDA streams (e.g. density evaluation and
        sampling requests served side by side): small-batch launches are then sized for num_sms / streams SMs so
        that the concurrent kernels are co-resident.  Every stream gets its own scratch workspace (`_workspace`)."""
        self.streams = max(1, int(streams))
        if self._plan is not None:
            _capi.check(_capi.lib().ipoke_inn_set_option(self._plan, b"streams", self.streams), "set_option")

    def invalidate(self):
        """Force a repack (and a fresh walk of the module's parameters) before the next call: after out-of-band
        ``param.data`` edits or after replacing Parameter objects inside the blocks."""
        self._packed_sig = None
        self._init_epoch = None
        self._tensors_cache = None
        self._grad_layout = None
        self._grad_ptr_offs = None

    def release(self):
        if self._plan is not None:
            _capi.lib().ipoke_inn_plan_destroy(self._plan)
        self._plan, self._packed, self._packed_sig, self._ws, self._device = None, None, None, {}, None
        self._pm = None                       # the Linear weights keep their (valid) storage inside the old master
        self._grad_layout = None
        self._grad_ptr_offs = None

    # copies / pickles of the module rebuild the plan, the packed buffer and the workspaces lazily (ctypes handles
    # cannot be pickled)
    def __getstate__(self):
        st = dict(self.__dict__)
        for k in ("_plan", "_packed", "_packed_sig", "_device", "_pm", "_init_epoch", "_tensors_cache", "_grad_layout",
                  "_grad_ptr_offs", "last_flat_grad", "grad_ready_hook", "_anchor"):
            st[k] = None
        st["_ws"] = {}
        return st

    def __setstate__(self, st):
        self.__dict__.update(st)

    def __deepcopy__(self, memo):
        import copy
        new = FlowEngine.__new__(FlowEngine)
        memo[id(self)] = new
        st = self.__getstate__()
        flow = st.pop("flow")
        desc = st.pop("desc")
        new.__dict__.update(copy.deepcopy(st, memo))
        new.flow = copy.deepcopy(flow, memo)
        new.desc = _capi.InnDesc.from_buffer_copy(desc)
        return new

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass

    def _tensors(self):
        # walked once and cached: the walk over 16 blocks x 4 subnetworks costs as much host time per call as a
        # third of the batch-64 forward.  Parameter objects survive .to() / load_state_dict() / optimizer steps; the
        # key (identity of every block's ActNorm parameters and of the first / last Linear) catches rebuilt blocks.
        blocks = self.flow.sub_layers
        key = tuple(id(b.norm_layer.loc) for b in blocks) + (
            id(blocks[0].coupling.s[0].linears()[0].weight), id(blocks[-1].coupling.t[1].linears()[-1].weight))
        cached = getattr(self, "_tensors_cache", None)
        if cached is not None and cached[0] == key:
            return cached[1]
        out = self._walk_tensors()
        self._tensors_cache = (key, out)
        self._grad_layout = None
        self._grad_ptr_offs = None
        return out

    def _walk_tensors(self):
        blocks = self.flow.sub_layers
        lin_w, lin_b = [], []
        for blk in blocks:
            for i in range(2):
                for nets in (blk.coupling.s, blk.coupling.t):
                    for lin in nets[i].linears():
                        lin_w.append(lin.weight)
                        lin_b.append(lin.bias)
        loc = [b.norm_layer.loc for b in blocks]
        scale = [b.norm_layer.scale for b in blocks]
        pf = [b.shuffle.forward_shuffle_idx for b in blocks]
        pb = [b.shuffle.backward_shuffle_idx for b in blocks]
        return lin_w, lin_b, loc, scale, pf, pb

    def _ensure(self, device):
        L = _capi.lib()
        if device.type != "cuda":
            raise RuntimeError("ipoke_inn_b200 runs on CUDA (sm_100a) only: move the module and its inputs to a "
                               "B200 device; there is no CPU fallback")
        if self._plan is None or self._device != device:
            self.release()
            plan = ctypes.c_void_p()
            idx = device.index if device.index is not None else torch.cuda.current_device()
            _capi.check(L.ipoke_inn_plan_create(ctypes.byref(self.desc), idx, ctypes.byref(plan)), "plan_create")
            self._plan = plan
            self._device = device
            self._packed = torch.empty(L.ipoke_inn_packed_bytes(plan), dtype=torch.uint8, device=device)
            self._packed_sig = None
            if self.streams > 1:
                _capi.check(L.ipoke_inn_set_option(plan, b"streams", self.streams), "set_option")
        return L

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self._device).cuda_stream)

    def _workspace(self, B):
        # one scratch workspace per (batch size, CUDA stream): calls issued on different streams (e.g. a density
        # evaluation and a sampling request served concurrently) must not share scratch state
        key = (B, torch.cuda.current_stream(self._device).cuda_stream)
        ws = self._ws.get(key)
        if ws is None:
            if len(self._ws) >= 8:
                self._ws.pop(next(iter(self._ws)))
            n = _capi.lib().ipoke_inn_workspace_bytes(self._plan, B, 1)
            ws = torch.empty(n, dtype=torch.uint8, device=self._device)
            self._ws[key] = ws
        return ws

    def _signature(self, tensors):
        lin_w, lin_b, loc, scale, pf, pb = tensors
        v = 0
        for group in (lin_w, lin_b, loc, scale, pf, pb):
            for t in group:
                v += t._version
        return (v, lin_w[0].data_ptr(), lin_w[-1].data_ptr(), loc[0].data_ptr())

    def pack(self, force=False):
        """Repack parameters into the kernel layout if they changed since the last pack."""
        if self.assume_frozen and self._packed_sig is not None and not force:
            return
        tensors = self._tensors()
        sig = self._signature(tensors)
        if not force and sig == self._packed_sig:
            return
        if self._pm is not None:
            if self._pm_valid(tensors):
                # packed-master mode: the Linear weights live in the fp32 master whose layout is the packed hi arena;
                # an out-of-band change (load_state_dict, ActNorm init, a foreign optimizer) re-splits the master and
                # refreshes the small part -- no gather / scatter pass
                pm = self._pm
                if pm.stale:
                    raise RuntimeError("ipoke_inn_b200: the fp32 master weights are sharded across ranks; call "
                                       "ShardedAdam.gather_master() before editing or repacking parameters")
                L = _capi.lib()
                _capi.check(L.ipoke_inn_master_to_packed(pm.master.data_ptr(), pm.hi.data_ptr(),
                                                         pm.lo.data_ptr() if pm.lo is not None else None, pm.n,
                                                         self._stream()), "master_to_packed")
                _capi.check(L.ipoke_inn_pack_small(self._plan, self._packed.data_ptr(), self._stream()), "pack_small")
                self._packed_sig = sig
                return
            self._pm = None                   # parameters were moved / replaced: back to the plain repack
            self._grad_layout = None
            self._grad_ptr_offs = None
        lin_w, lin_b, loc, scale, pf, pb = tensors
        for t in lin_w + lin_b + loc + scale:
            if t.dtype != torch.float32 or not t.is_contiguous() or t.device != self._device:
                raise RuntimeError("ipoke_inn_b200: parameters must be contiguous fp32 CUDA tensors on one device")
        for t in pf + pb:
            if t.dtype != torch.int64 or t.device != self._device:
                raise RuntimeError("ipoke_inn_b200: shuffle indices must be int64 tensors on the module's device")
        L = _capi.lib()
        _capi.check(L.ipoke_inn_pack(self._plan, _ptr_array(lin_w), _ptr_array(lin_b), _ptr_array(loc),
                                     _ptr_array(scale), _ptr_array(pf), _ptr_array(pb),
                                     ctypes.c_void_p(self._packed.data_ptr()), self._stream()), "pack")
        self._packed_sig = sig

    # ------------------------------------------------------------------ inputs
    def _prep(self, x, cond):
        if x.dtype != torch.float32:
            raise TypeError("ipoke_inn_b200 computes in fp32 I/O; got input dtype %s" % x.dtype)
        B = x.shape[0]
        x2 = x.reshape(B, -1)
        if x2.shape[1] != self.C:
            raise RuntimeError(f"expected {self.C} channels, got input of shape {tuple(x.shape)}")
        x2 = x2.contiguous()
        c2 = None
        if self.E > 0:
            if cond is None:
                raise RuntimeError("conditional flow needs a conditioning tensor")
            if cond.dtype != torch.float32 or cond.device != x.device:
                raise TypeError("conditioning must be an fp32 tensor on the input's device")
            c2 = cond.reshape(B, -1)
            if c2.shape[1] != self.E:
                raise RuntimeError(f"expected {self.E} conditioning channels, got {tuple(cond.shape)}")
            c2 = c2.contiguous()
        return B, x2, c2

    # ------------------------------------------------------------------ ActNorm data init
    def _maybe_actnorm_init(self, x2, c2):
        norms = [b.norm_layer for b in self.flow.sub_layers]
        epoch = sum(n._state_epoch for n in norms)
        if self._init_epoch == epoch:
            return
        flags = torch.stack([n.initialized for n in norms]).cpu()      # one host sync, first call only
        if bool((flags == 0).any()):
            self._actnorm_init(x2, c2, flags)
        self._init_epoch = epoch

    def _actnorm_init(self, x2, c2, flags):
        """Reference modules.py:255-275 + :288-290: every uninitialised ActNorm takes loc=-mean, scale=1/(std+1e-6)
        from its own input, block after block.  Moments are all-reduced across ranks when torch.distributed is
        initialised so that all replicas start identical."""
        L = _capi.lib()
        B = x2.shape[0]
        cur = x2
        dist = torch.distributed if (self.sync_actnorm_init and torch.distributed.is_available()
                                     and torch.distributed.is_initialized()) else None
        with torch.no_grad():
            for k, blk in enumerate(self.flow.sub_layers):
                if int(flags[k]) == 0:
                    mom = torch.zeros(2 * self.C + 1, dtype=torch.float64, device=self._device)
                    mom[-1] = B
                    _capi.check(L.ipoke_inn_actnorm_stats(cur.data_ptr(), B, self.C, mom.data_ptr(), self._stream()))
                    if dist is not None:
                        dist.all_reduce(mom)
                    count = float(mom[-1].item())
                    nl = blk.norm_layer
                    _capi.check(L.ipoke_inn_actnorm_finalize(mom.data_ptr(), count, self.C, nl.loc.data_ptr(),
                                                             nl.scale.data_ptr(), self._stream()))
                    nl.initialized.fill_(1)
                    self._packed_sig = None
                self.pack()
                nxt = torch.empty_like(cur)
                ld = torch.empty(B, dtype=torch.float32, device=self._device)
                self._forward_raw(cur, c2, nxt, ld, None, None, k, k + 1)
                cur = nxt
        self._packed_sig = None

    # ------------------------------------------------------------------ calls
    def _forward_raw(self, x2, c2, z, ld, outs, lds, kb, ke):
        L = _capi.lib()
        B = x2.shape[0]
        ws = self._workspace(B)
        _capi.check(L.ipoke_inn_forward(self._plan, self._packed.data_ptr(), x2.data_ptr(),
                                        c2.data_ptr() if c2 is not None else None, z.data_ptr(), ld.data_ptr(),
                                        outs.data_ptr() if outs is not None else None,
                                        lds.data_ptr() if lds is not None else None, B, kb, ke, ws.data_ptr(),
                                        self._stream()), "forward")
        self.launches = L.ipoke_inn_last_launch_count(self._plan)

    @annotate("inn.forward")
    def forward(self, x, cond, want_blocks=False):
        """-> z [B,C], logdet [B] (+ block_outs [n,B,C], block_logdets [n,B])"""
        B, x2, c2 = self._prep(x, cond)
        self._ensure(x2.device)
        self._maybe_actnorm_init(x2, c2)
        self.pack()
        z = torch.empty(B, self.C, dtype=torch.float32, device=self._device)
        ld = torch.empty(B, dtype=torch.float32, device=self._device)
        outs = lds = None
        if want_blocks:
            outs = torch.empty(self.n_flows, B, self.C, dtype=torch.float32, device=self._device)
            lds = torch.empty(self.n_flows, B, dtype=torch.float32, device=self._device)
        self._forward_raw(x2, c2, z, ld, outs, lds, 0, self.n_flows)
        return z, ld, outs, lds

    @annotate("inn.reverse")
    def reverse(self, z, cond):
        B, z2, c2 = self._prep(z, cond)
        L = self._ensure(z2.device)
        self.pack()
        x = torch.empty(B, self.C, dtype=torch.float32, device=self._device)
        ws = self._workspace(B)
        _capi.check(L.ipoke_inn_reverse(self._plan, self._packed.data_ptr(), z2.data_ptr(),
                                        c2.data_ptr() if c2 is not None else None, x.data_ptr(), B, 0, self.n_flows,
                                        ws.data_ptr(), self._stream()), "reverse")
        self.launches = L.ipoke_inn_last_launch_count(self._plan)
        return x

    # ------------------------------------------------------------------ sampling loop: several draws per conditioning row
    DRAW_ROW_TILE = 128        # ipoke_inn_*_draws needs the conditioning rows in multiples of one GEMM row tile

    def _prep_draws(self, x, cond, n_draws):
        """x: [n_draws, B, C] (any trailing singleton dims), cond: [B, E(,1,1)] -> padded 2-D views.
        B is padded to a multiple of 128 rows (zero latents / zero conditioning; the pad rows are dropped)."""
        if self.E <= 0:
            raise RuntimeError("multi-draw calls need a conditional flow")
        n_draws = int(n_draws)
        if x.dtype != torch.float32 or cond.dtype != torch.float32 or cond.device != x.device:
            raise TypeError("ipoke_inn_b200 computes in fp32 I/O on one device")
        if x.shape[0] != n_draws:
            raise RuntimeError(f"expected {n_draws} draws along dim 0, got {tuple(x.shape)}")
        B = cond.shape[0]
        c2 = cond.reshape(B, -1)
        x3 = x.reshape(n_draws, -1, self.C)
        if c2.shape[1] != self.E or x3.shape[1] != B:
            raise RuntimeError(f"expected draws [{n_draws},{B},{self.C}] and conditioning [{B},{self.E}], got "
                               f"{tuple(x.shape)} and {tuple(cond.shape)}")
        T = self.DRAW_ROW_TILE
        Bp = B if n_draws == 1 else (B + T - 1) // T * T
        if Bp != B:
            xp = x3.new_zeros(n_draws, Bp, self.C)
            xp[:, :B] = x3
            cp = c2.new_zeros(Bp, self.E)
            cp[:B] = c2
            x3, c2 = xp, cp
        return n_draws, B, Bp, x3.contiguous(), c2.contiguous()

    @annotate("inn.reverse_draws")
    def reverse_draws(self, z, cond, n_draws):
        """n_draws latents per conditioning row in one call: z [n_draws, B, C], cond [B, E] -> x [n_draws, B, C].
        The cond-only subnetworks of the 'cond' couplings run once per conditioning row (include/ipoke_inn.h)."""
        n, B, Bp, z3, c2 = self._prep_draws(z, cond, n_draws)
        L = self._ensure(z3.device)
        self.pack()
        x = torch.empty(n, Bp, self.C, dtype=torch.float32, device=self._device)
        ws = self._workspace(n * Bp)
        _capi.check(L.ipoke_inn_reverse_draws(self._plan, self._packed.data_ptr(), z3.data_ptr(), c2.data_ptr(),
                                              x.data_ptr(), Bp, n, ws.data_ptr(), self._stream()), "reverse_draws")
        self.launches = L.ipoke_inn_last_launch_count(self._plan)
        return x[:, :B]

    @annotate("inn.forward_draws")
    def forward_draws(self, x, cond, n_draws):
        """-> z [n_draws, B, C], logdet [n_draws, B] for n_draws inputs per conditioning row."""
        n, B, Bp, x3, c2 = self._prep_draws(x, cond, n_draws)
        L = self._ensure(x3.device)
        self._maybe_actnorm_init(x3.reshape(n * Bp, self.C)[:B], c2[:B])
        self.pack()
        z = torch.empty(n, Bp, self.C, dtype=torch.float32, device=self._device)
        ld = torch.empty(n, Bp, dtype=torch.float32, device=self._device)
        ws = self._workspace(n * Bp)
        _capi.check(L.ipoke_inn_forward_draws(self._plan, self._packed.data_ptr(), x3.data_ptr(), c2.data_ptr(),
                                              z.data_ptr(), ld.data_ptr(), Bp, n, ws.data_ptr(), self._stream()),
                    "forward_draws")
        self.launches = L.ipoke_inn_last_launch_count(self._plan)
        return z[:, :B], ld[:, :B]

    # ------------------------------------------------------------------ training (forward with saved state, backward)
    @annotate("inn.forward_train")
    def forward_train(self, x, cond):
        """Forward that keeps what the fused backward needs.  -> z, logdet, block_outs [n,B,C], block_logdets [n,B], saved"""
        B, x2, c2 = self._prep(x, cond)
        L = self._ensure(x2.device)
        self._maybe_actnorm_init(x2, c2)
        self.pack()
        dev = self._device
        z = torch.empty(B, self.C, dtype=torch.float32, device=dev)
        ld = torch.empty(B, dtype=torch.float32, device=dev)
        outs = torch.empty(self.n_flows, B, self.C, dtype=torch.float32, device=dev)
        lds = torch.empty(self.n_flows, B, dtype=torch.float32, device=dev)
        saved = torch.empty(L.ipoke_inn_saved_bytes(self._plan, B), dtype=torch.uint8, device=dev)
        ws = self._workspace(B)
        _capi.check(L.ipoke_inn_forward_train(self._plan, self._packed.data_ptr(), x2.data_ptr(),
                                              c2.data_ptr() if c2 is not None else None, z.data_ptr(), ld.data_ptr(),
                                              outs.data_ptr(), lds.data_ptr(), B, ws.data_ptr(), saved.data_ptr(),
                                              self._stream()), "forward_train")
        self.launches = L.ipoke_inn_last_launch_count(self._plan)
        return z, ld, outs, lds, saved, x2, c2

    def grad_anchor(self, device):
        """A 1-element leaf that requires grad: the autograd input standing in for the parameters in flat_grads_only mode."""
        a = getattr(self, "_anchor", None)
        if a is None or a.device != device:
            a = torch.zeros(1, device=device, requires_grad=True)
            self._anchor = a
        return a

    # ------------------------------------------------------------------ packed-master training (sharded optimizer)
    def _pm_valid(self, tensors=None):
        pm = self._pm
        if pm is None or self._packed is None or pm.packed_ptr != self._packed.data_ptr():
            return False
        t = tensors or self._tensors()
        lin_w, lin_b = t[0], t[1]
        base = pm.master.data_ptr()
        return (lin_w[0].data_ptr() == base + 4 * pm.w_off[0] and lin_w[-1].data_ptr() == base + 4 * pm.w_off[-1]
                and lin_b[0].data_ptr() == base + 4 * pm.n)

    def adopt_packed_master(self):
        """Make one flat fp32 buffer whose element layout IS the packed hi arena the storage of every nn.Linear weight
        (``weight.data`` becomes a view of it; names, shapes, values and the Parameter objects are unchanged).  An
        optimizer shard can then own a contiguous span of that buffer and write the fp32 master and the bf16 hi / lo
        images the GEMMs read in one pass, and an all-gather can deliver the other ranks' spans straight into the
        packed buffer (optim_sharded.ShardedAdam).  Needs shapes without padding (H, E, C/2 multiples of 64: every
        reference config); raises NotImplementedError otherwise.  -> the packed-master record."""
        import types
        lin_w = self._tensors()[0]
        dev = lin_w[0].device
        L = self._ensure(dev)
        if self._pm is not None and self._pm_valid():
            return self._pm
        n_lin = len(lin_w)
        n = 8 + n_lin + 2 * self.n_flows
        out = (ctypes.c_longlong * n)()
        _capi.check(L.ipoke_inn_param_layout(self._plan, out, n), "param_layout")
        if out[0] != 1:
            raise NotImplementedError("packed-master training needs H, E and C/2 to be multiples of 64 (no padding inside "
                                      "the packed weight arena); use FusedAdam + GradSync for this shape")
        self._pm = None
        N = int(out[2])
        _, lin_b, loc, scale, _, _ = self._tensors()
        small = list(lin_b) + list(loc) + list(scale)       # the order of the flat gradient buffer behind the weights
        n_small = sum(p.numel() for p in small)
        master = torch.zeros(N + n_small, dtype=torch.float32, device=dev)
        w_off = [int(out[8 + i]) for i in range(n_lin)]
        with torch.no_grad():
            off = N
            for p in small:                                  # biases / ActNorm: one flat span behind the weights
                v = master[off:off + p.numel()].view(p.shape)
                v.copy_(p)
                p.data = v
                off += p.numel()
            self._packed_sig = None
            self.pack(force=True)                            # plain repack: its job tables keep the small parameters'
            for w, off in zip(lin_w, w_off):                 # (new) addresses; the weights still own their storage here
                v = master[off:off + w.numel()].view(w.shape)
                v.copy_(w)
                w.data = v
        pk16 = self._packed.view(torch.bfloat16)
        lo_off = int(out[3])
        self._pm = types.SimpleNamespace(
            master=master, w_off=w_off, n=N, n_small=n_small, packed_ptr=self._packed.data_ptr(), stale=False,
            hi=pk16[:N], lo=pk16[lo_off // 2:lo_off // 2 + N] if lo_off >= 0 else None,
            block_ranges=[(int(out[8 + n_lin + 2 * k]), int(out[8 + n_lin + 2 * k + 1])) for k in range(self.n_flows)])
        self._grad_layout = None
        self._grad_ptr_offs = None
        self._packed_sig = self._signature(self._tensors())  # the packed images equal the master just copied
        return self._pm

    def grad_layout(self):
        """Flat fp32 gradient buffer layout.  Plain mode: every parameter of the flow in ``flow.parameters()`` order
        (block by block), so a block's gradients are one contiguous slice (per-block all-reduce chunks).  Packed-master
        mode: the Linear weights first, in the element order of the packed hi arena (block by block as well), then the
        biases and the ActNorm parameters.  -> (offsets by id, total, per-block ranges, params in parameters() order)."""
        if getattr(self, "_grad_layout", None) is None and self._pm is not None:
            lin_w, lin_b, loc, scale, _, _ = self._tensors()
            pm = self._pm
            order = sorted(range(len(lin_w)), key=lambda i: pm.w_off[i])
            flat_params = [lin_w[i] for i in order] + list(lin_b) + list(loc) + list(scale)
            offs, cur = {}, 0
            for p in flat_params:
                offs[id(p)] = (cur, p.numel(), tuple(p.shape))
                cur += p.numel()
            for i, w in enumerate(lin_w):
                assert offs[id(w)][0] == pm.w_off[i], "packed hi arena is not the concatenation of the Linear weights"
            params = [p for blk in self.flow.sub_layers for p in blk.parameters()]
            pos = {id(p): j for j, p in enumerate(flat_params)}
            self._grad_flat_params = flat_params
            self._grad_perm = [pos[id(p)] for p in params]
            self._grad_layout = (offs, cur, list(pm.block_ranges), params)
        if getattr(self, "_grad_layout", None) is None:
            self._grad_flat_params = None
            self._grad_perm = None
            offs, cur = {}, 0
            block_ranges = []
            for k, blk in enumerate(self.flow.sub_layers):
                start = cur
                for p in blk.parameters():
                    offs[id(p)] = (cur, p.numel(), tuple(p.shape))
                    cur += p.numel()
                block_ranges.append((start, cur))
            params = [p for blk in self.flow.sub_layers for p in blk.parameters()]
            self._grad_layout = (offs, cur, block_ranges, params)
        return self._grad_layout

    def new_grad_buffer(self, want_views=True):
        """-> flat [total] fp32 (fresh storage: gradients handed to autograd must not alias a later call's),
        the per-parameter views in ``flow.parameters()`` order (None when not wanted) and the pointer arrays of the C ABI."""
        offs, total, block_ranges, params = self.grad_layout()
        flat = torch.empty(total, dtype=torch.float32, device=self._device)
        base = flat.data_ptr()
        if getattr(self, "_grad_ptr_offs", None) is None:          # byte offsets per C-ABI pointer array, built once
            lin_w, lin_b, loc, scale, _, _ = self._tensors()
            self._grad_ptr_offs = [[4 * offs[id(t)][0] for t in ts] for ts in (lin_w, lin_b, loc, scale)]
        arrays = tuple((ctypes.c_void_p * len(o))(*[base + b for b in o]) for o in self._grad_ptr_offs)
        if not want_views:
            return flat, None, arrays
        # one C++ call for all the views (a Python slice + view per parameter costs more than the batch-64 kernels)
        if self._grad_perm is None:
            views = list(torch._utils._unflatten_dense_tensors(flat, params))
        else:
            fv = torch._utils._unflatten_dense_tensors(flat, self._grad_flat_params)
            views = [fv[j] for j in self._grad_perm]
        return flat, views, arrays

    @annotate("inn.backward")
    def backward(self, x2, c2, outs, saved, dz, dld, grad_ptrs, want_dx=True, want_dcond=True, block_begin=0,
                 block_end=None, dcond=None):
        """Fused backward over blocks [block_begin, block_end).  -> dx [B,C] or None, dcond [B,E] or None."""
        L = _capi.lib()
        B = x2.shape[0]
        block_end = self.n_flows if block_end is None else block_end
        dev = self._device
        dx = torch.empty(B, self.C, dtype=torch.float32, device=dev) if want_dx else None
        if self.E > 0 and want_dcond and dcond is None:
            dcond = torch.empty(B, self.E, dtype=torch.float32, device=dev)
        gw, gb, gloc, gscale = grad_ptrs
        ws = self._workspace(B)
        _capi.check(L.ipoke_inn_backward(self._plan, self._packed.data_ptr(), x2.data_ptr(),
                                         c2.data_ptr() if c2 is not None else None, outs.data_ptr(), saved.data_ptr(),
                                         dz.data_ptr(), dld.data_ptr(), dx.data_ptr() if dx is not None else None,
                                         dcond.data_ptr() if dcond is not None else None, gw, gb, gloc, gscale,
                                         B, block_begin, block_end, ws.data_ptr(), self._stream()), "backward")
        self.bwd_launches = L.ipoke_inn_last_launch_count(self._plan)
        return dx, dcond

    # host-buffer entry points (pinned CPU tensors in, pinned CPU tensors out; stream-ordered, caller syncs)
    def _ready_for_host_call(self):
        if self._plan is None or self._packed is None:
            raise RuntimeError("ipoke_inn_b200: run one device call first (it creates the plan on the module's device "
                               "and does the ActNorm data init), then use the *_host entry points")
        self.pack()

    @annotate("inn.forward_host")
    def forward_host(self, x_host, cond_host, z_host, logdet_host):
        L = _capi.lib()
        self._ready_for_host_call()
        B = x_host.shape[0]
        ws = self._workspace(B)
        _capi.check(L.ipoke_inn_forward_host(self._plan, self._packed.data_ptr(), x_host.data_ptr(),
                                             cond_host.data_ptr() if cond_host is not None else None,
                                             z_host.data_ptr(), logdet_host.data_ptr(), B, ws.data_ptr(),
                                             self._stream()), "forward_host")
        self.launches = L.ipoke_inn_last_launch_count(self._plan)

    @annotate("inn.reverse_host")
    def reverse_host(self, z_host, cond_host, x_host):
        L = _capi.lib()
        self._ready_for_host_call()
        B = z_host.shape[0]
        ws = self._workspace(B)
        _capi.check(L.ipoke_inn_reverse_host(self._plan, self._packed.data_ptr(), z_host.data_ptr(),
                                             cond_host.data_ptr() if cond_host is not None else None,
                                             x_host.data_ptr(), B, ws.data_ptr(), self._stream()), "reverse_host")
        self.launches = L.ipoke_inn_last_launch_count(self._plan)
