"""Training path: one autograd.Function per flow call over the fused forward (with saved state) and the fused
backward kernels, so that the caller's ``loss.backward()`` (Lightning's loop in the reference,
models/second_stage_video_fc.py:448-490) reaches ``ipoke_inn_backward``.

Parameter gradients are written by the kernels straight into ONE flat fp32 buffer laid out in
``flow.parameters()`` order; autograd receives views of it (AccumulateGrad adopts them without a copy when
``.grad`` is None), and ``engine.last_flat_grad`` / ``engine.grad_layout()`` expose the buffer so that the
gradient all-reduce of the training step is a few large NCCL calls on contiguous memory (parallel.GradSync).
``engine.direct_param_grads = True`` (opt-in) assigns the views to ``p.grad`` inside backward() instead.
"""
import torch


class FlowFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, engine, x, cond, *params):
        z, ld, outs, lds, saved, x2, c2 = engine.forward_train(x, cond)
        ctx.engine = engine
        ctx.saved_buf = saved
        ctx.x_shape = x.shape
        ctx.cond_shape = None if cond is None else cond.shape
        ctx.save_for_backward(x2, c2, outs)
        ctx.mark_non_differentiable(outs, lds)
        return z, ld, outs, lds

    @staticmethod
    def backward(ctx, dz, dld, _douts, _dlds):
        engine = ctx.engine
        x2, c2, outs = ctx.saved_tensors
        need_x = ctx.needs_input_grad[1]
        need_c = c2 is not None and ctx.needs_input_grad[2]
        flat_only = getattr(engine, "flat_grads_only", False)
        flat, views, ptrs = engine.new_grad_buffer() if not flat_only else engine.new_grad_buffer(want_views=False)
        dz = dz.contiguous().float()
        dld = dld.contiguous().float()
        hook = engine.grad_ready_hook
        if hook is None:
            dx, dcond = engine.backward(x2, c2, outs, ctx.saved_buf, dz, dld, ptrs, want_dx=need_x, want_dcond=need_c)
        else:
            # chunked by flow block (last first) so that the hook can start reducing a block's gradients while
            # the earlier blocks are still being computed
            n = engine.n_flows
            chunk = max(1, int(engine.grad_chunk_blocks))
            dcond = None
            g = dz
            ke = n
            while ke > 0:
                kb = max(0, ke - chunk)
                last = kb == 0
                dx, dcond = engine.backward(x2, c2, outs, ctx.saved_buf, g, dld, ptrs, want_dx=(not last) or need_x,
                                            want_dcond=need_c, block_begin=kb, block_end=ke, dcond=dcond)
                hook(engine, flat, kb, ke)
                g = dx
                ke = kb
        engine.last_flat_grad = flat
        # ctx.saved_buf stays alive with the graph (retain_graph=True / a second backward read it again)
        gx = dx.view(ctx.x_shape) if (need_x and dx is not None) else None
        gc = dcond.view(ctx.cond_shape) if (need_c and dcond is not None) else None
        if flat_only:
            # the optimizer reads engine.last_flat_grad itself (optim_sharded.ShardedAdam): no per-parameter views, no
            # AccumulateGrad work -- at the reference's batch of 64 that bookkeeping is a third of the step's host time
            return (None, gx, gc) + (None,) * (len(ctx.needs_input_grad) - 3)
        if engine.direct_param_grads:
            # opt-in (engine.direct_param_grads = True): hand the views to the parameters here instead of through one
            # AccumulateGrad node per parameter -- the autograd engine's per-tensor bookkeeping for the few hundred
            # parameters of a flow costs more host time than the batch-64 kernels take.  Same values in ``.grad``
            # (adopted when it is None, accumulated otherwise); per-parameter autograd hooks do not fire and
            # ``torch.autograd.grad(..., params)`` sees None, which is why it is not the default.
            params = engine.grad_layout()[3]
            acc_p, acc_v = [], []
            for p, v in zip(params, views):
                if p.grad is None:
                    p.grad = v
                else:
                    acc_p.append(p.grad)
                    acc_v.append(v)
            if acc_p:
                torch._foreach_add_(acc_p, acc_v)
            return (None, gx, gc) + (None,) * len(views)
        return (None, gx, gc) + tuple(views)
