"""Sharded optimizer step of the data-parallel training step (BASELINE config 4).

The reference trains with Lightning DDP: a bucketed NCCL all-reduce of every gradient (950.8 MB of fp32 at cfg A,
experiments/experiment.py:86) followed by a replicated ``torch.optim.Adam(lr, betas, weight_decay, amsgrad=True)``
(models/second_stage_video_fc.py:670-700).  ``ShardedAdam`` keeps those semantics (mean gradient over ranks, Adam /
AMSGrad with L2 weight decay, arithmetic of ``ipoke_inn_adam_step``) and changes where the bytes go:

* the nn.Linear weights (99.9 % of the parameters) live in one flat fp32 *master* whose element layout is the packed hi
  arena of the kernels (``FlowEngine.adopt_packed_master``), and the fused backward writes their gradients into a flat
  buffer of the same layout;
* the backward pass is issued in chunks of flow blocks (last block first); as soon as a chunk's kernels are enqueued,
  its contiguous gradient span is **reduce-scattered** on a side stream (fp32, or bf16 on the wire), overlapping the
  backward of the earlier blocks;
* every rank runs ONE Adam pass over the 1/N span it owns, which writes the fp32 master, the moments (allocated for
  the owned span only) and the bf16 hi / lo images of the weights straight into the packed buffer;
* the bf16 images are **all-gathered in place inside the packed buffer** -- the wire carries 2 (bf16 mode) or 4
  (parity mode) bytes per weight instead of fp32 gradients a second time, the 36-byte-per-parameter optimizer traffic
  drops to 1/N, and the separate fp32 -> bf16 repack pass of the replicated path disappears (also at N = 1);
* biases and ActNorm parameters (0.1 % of the bytes) sit in one flat span behind the weights; their gradients are
  all-reduced in one small message and every rank updates them replicated with the same kernel;
* the backward pass hands the optimizer ONE flat gradient buffer and creates no per-parameter ``.grad`` tensors
  (``expose_param_grads=True`` restores them): at the reference's batch of 64 the views and AccumulateGrad bookkeeping
  of 336 parameters were a third of the step's host time.

The fp32 master of the spans a rank does not own goes stale between checkpoints: ``gather_master()`` (called
automatically by the module's ``state_dict()``) all-gathers it.
"""
import torch

from ._nvtx import annotate
import torch.distributed as dist

from . import _capi


def plan_pieces(block_ranges, chunk_blocks, world, rank):
    """Backward chunks (last flow block first) and the span of each chunk this rank owns.
    -> list of dicts {kb, ke, lo, hi, own_lo, own_n, sharded}; a chunk is sharded when its length divides by
    4 * world (float4 kernels, equal NCCL counts), else every rank owns all of it (all-reduce + replicated update)."""
    n = len(block_ranges)
    chunk_blocks = max(1, int(chunk_blocks))
    out, ke = [], n
    while ke > 0:
        kb = max(0, ke - chunk_blocks)
        lo, hi = block_ranges[kb][0], block_ranges[ke - 1][1]
        length = hi - lo
        sharded = world > 1 and length % (4 * world) == 0
        if sharded:
            own_n = length // world
            own_lo = lo + rank * own_n
        else:
            own_n, own_lo = length, lo
        out.append(dict(kb=kb, ke=ke, lo=lo, hi=hi, own_lo=own_lo, own_n=own_n, sharded=sharded))
        ke = kb
    return out


class ShardedAdam(torch.optim.Optimizer):
    """Adam / AMSGrad for one flow stack with reduce-scatter -> shard update -> all-gather into the packed weights.

        net = SupervisedTransformer(cfg).cuda()
        opt = ShardedAdam(net, lr=1e-3, weight_decay=1e-5, amsgrad=True)     # after .cuda(); all ranks, same seed
        loss.backward(); opt.step()

    ``param_groups[0]["lr"]`` may be rewritten every step (the reference's warm-up, second_stage_video_fc.py:256-271).
    ``wire="bf16"`` sends the gradients as bf16 (pre-divided by the world size; the sum on the wire is the mean).
    One backward pass per step (no gradient accumulation across backward calls).  Parameters' ``.grad`` stay None unless
    ``expose_param_grads=True``; ``gradients()`` returns views of this rank's (un-reduced) flat gradient buffer on demand.
    """

    def __init__(self, net, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, amsgrad=False, group=None,
                 wire="fp32", chunk_blocks=2, expose_param_grads=False):
        if wire not in ("fp32", "bf16"):
            raise ValueError("wire must be 'fp32' or 'bf16'")
        if lr < 0.0 or eps < 0.0 or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0 or weight_decay < 0.0:
            raise ValueError("ShardedAdam: invalid hyper-parameter")
        flow = getattr(net, "flow", net)
        self.engine = flow.engine
        params = [p for blk in flow.sub_layers for p in blk.parameters()]
        if not params or params[0].device.type != "cuda":
            raise RuntimeError("ShardedAdam: move the module to its CUDA device before building the optimizer")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad))
        self.group = group
        self.wire = wire
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        eng = self.engine
        pm = eng.adopt_packed_master()
        self.pm = pm
        self.device = pm.master.device
        self.pieces = plan_pieces(pm.block_ranges, chunk_blocks, self.world, self.rank)
        self._by_range = {(c["kb"], c["ke"]): i for i, c in enumerate(self.pieces)}
        own = sum(c["own_n"] for c in self.pieces)
        self.m = torch.zeros(own, dtype=torch.float32, device=self.device)
        self.v = torch.zeros(own, dtype=torch.float32, device=self.device)
        self.vmax = torch.zeros(own, dtype=torch.float32, device=self.device) if amsgrad else None
        off = 0
        for c in self.pieces:
            c["m_off"] = off
            off += c["own_n"]
            if c["sharded"]:
                c["g"] = torch.empty(c["own_n"], dtype=torch.bfloat16 if wire == "bf16" else torch.float32,
                                     device=self.device)
        self._stage16 = (torch.empty(pm.n, dtype=torch.bfloat16, device=self.device)
                         if (wire == "bf16" and self.world > 1) else None)
        # biases / ActNorm: replicated moments for the flat span behind the weights
        self.sm = torch.zeros(pm.n_small, dtype=torch.float32, device=self.device)
        self.sv = torch.zeros(pm.n_small, dtype=torch.float32, device=self.device)
        self.svmax = torch.zeros(pm.n_small, dtype=torch.float32, device=self.device) if amsgrad else None
        eng.flat_grads_only = not expose_param_grads
        self.steps = 0
        self._side = None
        self._reduced = None                  # the flat gradient buffer whose chunks are in flight / reduced
        if self.world > 1:
            eng.grad_chunk_blocks = chunk_blocks
            eng.grad_ready_hook = self
        self._sd_hook = net.register_state_dict_pre_hook(lambda *a, **k: self.gather_master()) \
            if hasattr(net, "register_state_dict_pre_hook") else None

    # ------------------------------------------------------------------ gradient exchange (called from the backward)
    def _side_stream(self):
        if self._side is None:
            self._side = torch.cuda.Stream(device=self.device)
        return self._side

    def __call__(self, engine, flat, block_begin, block_end):
        """engine.grad_ready_hook: the kernels of blocks [block_begin, block_end) are enqueued on the current stream."""
        c = self.pieces[self._by_range[(block_begin, block_end)]]
        cur = torch.cuda.current_stream(self.device)
        side = self._side_stream()
        side.wait_stream(cur)
        span = flat[c["lo"]:c["hi"]]
        L = _capi.lib()
        with torch.cuda.stream(side):
            if not c["sharded"]:
                dist.all_reduce(span, op=dist.ReduceOp.AVG, group=self.group)
            elif self.wire == "fp32":
                dist.reduce_scatter_tensor(c["g"], span, op=dist.ReduceOp.AVG, group=self.group)
            else:
                st16 = self._stage16[c["lo"]:c["hi"]]
                _capi.check(L.ipoke_inn_grad_to_bf16(span.data_ptr(), st16.data_ptr(), span.numel(), 1.0 / self.world,
                                                     side.cuda_stream), "grad_to_bf16")
                dist.reduce_scatter_tensor(c["g"], st16, op=dist.ReduceOp.SUM, group=self.group)
            if block_begin == 0:                                        # biases + ActNorm parameters: one small message
                dist.all_reduce(flat[self.pm.n:], op=dist.ReduceOp.AVG, group=self.group)
        flat.record_stream(side)
        self._reduced = flat

    # ------------------------------------------------------------------ the step
    @torch.no_grad()
    @annotate("optim.step")
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        eng, pm = self.engine, self.pm
        if eng._pm is not pm or not eng._pm_valid():
            raise RuntimeError("ShardedAdam: the module's parameters were moved or replaced after the optimizer was built")
        flat = eng.last_flat_grad
        if flat is None:
            raise RuntimeError("ShardedAdam.step() needs the gradients of one backward pass through the flow")
        if self.world > 1 and self._reduced is not flat:
            raise RuntimeError("ShardedAdam: the gradient exchange did not run for this backward pass (was "
                               "engine.grad_ready_hook replaced?)")
        g = self.param_groups[0]
        lr, (b1, b2), eps, wd, ams = float(g["lr"]), g["betas"], float(g["eps"]), float(g["weight_decay"]), bool(g["amsgrad"])
        self.steps += 1
        L = _capi.lib()
        cur = torch.cuda.current_stream(self.device)
        side = self._side_stream() if self.world > 1 else None
        if side is not None:
            cur.wait_stream(side)                                       # reductions done (device-side wait only)
        for c in self.pieces:
            lo, n, mo = c["own_lo"], c["own_n"], c["m_off"]
            g32 = g16 = None
            if c["sharded"]:
                if self.wire == "bf16":
                    g16 = c["g"].data_ptr()
                else:
                    g32 = c["g"].data_ptr()
            else:
                g32 = flat.data_ptr() + 4 * lo
            _capi.check(L.ipoke_inn_adam_shard_step(
                pm.master.data_ptr() + 4 * lo, g32, g16, self.m.data_ptr() + 4 * mo, self.v.data_ptr() + 4 * mo,
                (self.vmax.data_ptr() + 4 * mo) if self.vmax is not None else None,
                pm.hi.data_ptr() + 2 * lo, (pm.lo.data_ptr() + 2 * lo) if pm.lo is not None else None, n,
                lr, float(b1), float(b2), eps, wd, self.steps, int(ams), 1.0, cur.cuda_stream), "adam_shard_step")
            if c["sharded"]:
                # the owned span's new bf16 images -> every rank's packed buffer, in place, while the next chunk updates
                side.wait_stream(cur)
                with torch.cuda.stream(side):
                    dist.all_gather_into_tensor(pm.hi[c["lo"]:c["hi"]], pm.hi[lo:lo + n], group=self.group)
                    if pm.lo is not None:
                        dist.all_gather_into_tensor(pm.lo[c["lo"]:c["hi"]], pm.lo[lo:lo + n], group=self.group)
        # biases / ActNorm: replicated (their gradients were all-reduced with the last chunk), same kernel, no images
        _capi.check(L.ipoke_inn_adam_shard_step(
            pm.master.data_ptr() + 4 * pm.n, flat.data_ptr() + 4 * pm.n, None, self.sm.data_ptr(), self.sv.data_ptr(),
            self.svmax.data_ptr() if self.svmax is not None else None, None, None, pm.n_small,
            lr, float(b1), float(b2), eps, wd, self.steps, int(ams), 1.0, cur.cuda_stream), "adam_shard_step(small)")
        _capi.check(L.ipoke_inn_pack_small(eng._plan, eng._packed.data_ptr(), cur.cuda_stream), "pack_small")
        if side is not None:
            cur.wait_stream(side)
        pm.stale = any(c["sharded"] for c in self.pieces)
        t = eng._tensors()
        torch.autograd.graph.increment_version(list(t[0]) + list(t[1]) + list(t[2]) + list(t[3]))
        eng._packed_sig = eng._signature(t)                            # the packed buffer is current: no repack
        eng.last_flat_grad = None
        self._reduced = None
        return loss

    # ------------------------------------------------------------------ fp32 master on every rank (checkpoints)
    @torch.no_grad()
    def gather_master(self):
        """All-gather the fp32 master spans so that every rank's parameters are current.  A COLLECTIVE: every rank must
        call it (and therefore ``net.state_dict()``, which triggers it) at the same point; a no-op when nothing is stale."""
        pm = self.pm
        if not pm.stale:
            return
        for c in self.pieces:
            if c["sharded"]:
                lo, n = c["own_lo"], c["own_n"]
                dist.all_gather_into_tensor(pm.master[c["lo"]:c["hi"]], pm.master[lo:lo + n], group=self.group)
        pm.stale = False

    def gradients(self):
        """Per-parameter views (``flow.parameters()`` order) of the flat gradient buffer of the last backward pass: this
        rank's own gradients -- the mean over ranks exists only as the reduce-scattered spans -- until ``step()``."""
        eng = self.engine
        flat = eng.last_flat_grad
        if flat is None:
            raise RuntimeError("no gradient buffer: run a backward pass first (step() consumes it)")
        offs, _, _, params = eng.grad_layout()
        return [flat[offs[id(p)][0]:offs[id(p)][0] + p.numel()].view(p.shape) for p in params]

    def state_dict(self):
        """This rank's optimizer shard (moments of the owned spans) plus the replicated small-parameter moments."""
        return {"steps": self.steps, "rank": self.rank, "world": self.world, "wire": self.wire,
                "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups],
                "m": self.m, "v": self.v, "vmax": self.vmax, "sm": self.sm, "sv": self.sv, "svmax": self.svmax}

    def load_state_dict(self, sd):
        if sd["world"] != self.world or sd["rank"] != self.rank or sd["m"].numel() != self.m.numel():
            raise RuntimeError("ShardedAdam.load_state_dict: shard layout differs (world size / rank / architecture)")
        self.steps = int(sd["steps"])
        self.m.copy_(sd["m"])
        self.v.copy_(sd["v"])
        if self.vmax is not None and sd.get("vmax") is not None:
            self.vmax.copy_(sd["vmax"])
        for g, s in zip(self.param_groups, sd["param_groups"]):
            g.update(s)
        self.sm.copy_(sd["sm"])
        self.sv.copy_(sd["sv"])
        if self.svmax is not None and sd.get("svmax") is not None:
            self.svmax.copy_(sd["svmax"])

    def remove(self):
        if self.engine.grad_ready_hook is self:
            self.engine.grad_ready_hook = None
        self.engine.flat_grads_only = False
        if self._sd_hook is not None:
            self._sd_hook.remove()
            self._sd_hook = None
