"""Data parallelism of the INN path: one process per GPU, the batch sharded by rank, weights replicated.

Forward and reverse need no communication (samples are independent).  The training
step has ONE exchange: the sum of the INN gradients over ranks -- what Lightning's implicit DDP does for the
reference (experiments/experiment.py:86, bucketed NCCL all-reduce).  Here the fused backward writes every
gradient into one flat fp32 buffer in ``flow.parameters()`` order and is issued in chunks of flow blocks
(last block first); ``GradSync`` all-reduces each chunk's contiguous slice on a side stream as soon as its
kernels are enqueued, so the reduction of blocks 7,6 overlaps the backward of blocks 5..0, and the compute
stream waits for the side stream (an event, no host sync) before ``loss.backward()`` returns.

ActNorm's data-dependent initialisation is made rank-consistent by all-reducing its moments
(engine.FlowEngine._actnorm_init); Shuffle permutations are identical when every rank seeds identically,
as the reference does (experiments/experiment.py:29-36) -- ``broadcast_parameters`` makes it explicit.
"""
import torch
import torch.distributed as dist


def shard_bounds(n, rank, world):
    """Contiguous shard [lo, hi) of n samples for `rank` (sizes differ by at most one)."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_batch(t, rank=None, world=None):
    if t is None:
        return None
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    lo, hi = shard_bounds(t.shape[0], rank, world)
    return t[lo:hi]


def broadcast_parameters(module, src=0, group=None):
    """Make parameters and buffers (ActNorm state, Shuffle indices) identical on every rank."""
    with torch.no_grad():
        for t in list(module.parameters()) + list(module.buffers()):
            dist.broadcast(t.data, src=src, group=group)
    flow = getattr(module, "flow", module)
    eng = getattr(flow, "engine", None)
    if eng is not None:
        eng.invalidate()
    elif hasattr(flow, "invalidate"):          # MaCow flows: writes through .data do not bump the version counters
        flow.invalidate()                      # their packed segments are keyed on


class GradSync:
    """Gradient reduction of one flow, overlapped with its backward pass.

        sync = GradSync(net.flow)            # installs engine.grad_ready_hook
        loss.backward()                      # chunks are reduced on a side stream while earlier blocks compute
        optimizer.step()                     # gradients are already the mean over ranks (DDP semantics)

    average=True divides by the world size (DDP); chunk_blocks = flow blocks per backward chunk / message.
    """

    def __init__(self, flow, group=None, average=True, chunk_blocks=2):
        self.flow = getattr(flow, "flow", flow)
        self.engine = self.flow.engine
        self.group = group
        self.average = average
        self.engine.grad_chunk_blocks = chunk_blocks
        self.engine.grad_ready_hook = self
        self._stream = None
        self._pending = []
        self.messages = []          # (element offset, element count) of the last backward, for tests / accounting

    def remove(self):
        self.engine.grad_ready_hook = None

    def _world(self):
        return dist.get_world_size(self.group) if dist.is_available() and dist.is_initialized() else 1

    def __call__(self, engine, flat, block_begin, block_end):
        """Called by backward.FlowFunction right after the kernels of blocks [block_begin, block_end) are enqueued."""
        ranges = engine.grad_layout()[2]
        lo, hi = ranges[block_begin][0], ranges[block_end - 1][1]
        if block_end == engine.n_flows:
            self.messages = []
        self.messages.append((lo, hi - lo))
        world = self._world()
        if world > 1:
            chunk = flat[lo:hi]
            if flat.is_cuda:
                cur = torch.cuda.current_stream(flat.device)
                if self._stream is None:
                    self._stream = torch.cuda.Stream(device=flat.device)
                self._stream.wait_stream(cur)                     # the chunk's kernels precede the reduction
                with torch.cuda.stream(self._stream):
                    self._reduce(chunk, world)
                chunk.record_stream(self._stream)
            else:
                self._reduce(chunk, world)
        if block_begin == 0:
            if getattr(engine, "_pm", None) is not None and world > 1:
                # packed-master layout: biases and ActNorm parameters sit behind the weight arena, not inside the
                # per-block ranges -- one more (small) message
                lo, hi = engine._pm.n, flat.numel()
                self.messages.append((lo, hi - lo))
                tail = flat[lo:hi]
                if flat.is_cuda:
                    self._stream.wait_stream(torch.cuda.current_stream(flat.device))
                    with torch.cuda.stream(self._stream):
                        self._reduce(tail, world)
                else:
                    self._reduce(tail, world)
            self.finish(flat)

    def _reduce(self, chunk, world):
        backend = dist.get_backend(self.group)
        if self.average and backend == "nccl":
            dist.all_reduce(chunk, op=dist.ReduceOp.AVG, group=self.group)
        else:
            dist.all_reduce(chunk, op=dist.ReduceOp.SUM, group=self.group)
            if self.average:
                chunk.mul_(1.0 / world)

    def finish(self, flat=None):
        """Order the compute stream after the outstanding reductions (device-side wait only)."""
        if self._stream is not None and flat is not None and flat.is_cuda:
            torch.cuda.current_stream(flat.device).wait_stream(self._stream)
