"""Drop-in for the reference's ``scene/OurAdam.py`` (class ``Adam``, :8-175): an Adam optimizer whose
``step(relevant)`` only updates the rows listed in ``relevant`` (moments of other rows stay stale,
no decay), and updates every row when ``relevant`` is empty.

Same constructor, ``param_groups`` and ``state`` layout (``'step'`` CPU scalar tensor, ``'exp_avg'``,
``'exp_avg_sq'``), so ``state_dict()`` checkpoints are interchangeable.  The arithmetic of
``_single_tensor_adam`` / ``_single_tensor_adam2`` (:249-426) runs as one fused CUDA launch over all
parameter tensors of the step (h3dgs_sparse_adam, include/h3dgs_b200.h) instead of ~10 torch kernels
and three gathered temporaries per parameter group.
"""
import ctypes
import math
from typing import Optional

import torch
from torch.optim.optimizer import Optimizer

import h3dgs_native as nat


class Adam(Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, *,
                 foreach: Optional[bool] = None, maximize: bool = False, capturable: bool = False):
        if not 0.0 <= lr:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if not 0.0 <= eps:
            raise ValueError("Invalid epsilon value: {}".format(eps))
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError("Invalid beta parameter at index 0: {}".format(betas[0]))
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError("Invalid beta parameter at index 1: {}".format(betas[1]))
        if not 0.0 <= weight_decay:
            raise ValueError("Invalid weight_decay value: {}".format(weight_decay))
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad,
                        maximize=maximize, foreach=foreach, capturable=capturable)
        super(Adam, self).__init__(params, defaults)

    def __setstate__(self, state):
        super().__setstate__(state)
        for group in self.param_groups:
            group.setdefault('amsgrad', False)
            group.setdefault('maximize', False)
            group.setdefault('foreach', None)
            group.setdefault('capturable', False)
        state_values = list(self.state.values())
        step_is_tensor = (len(state_values) != 0) and torch.is_tensor(state_values[0]['step'])
        if not step_is_tensor:
            for s in state_values:
                s['step'] = torch.tensor(float(s['step']))

    def _update_terms(self, group, p):
        """Advance the parameter's step count and return the host-side terms of the update
        (scene/OurAdam.py:309-325: bias corrections in double precision on the host)."""
        beta1, beta2 = group['betas']
        state = self.state[p]
        if len(state) == 0:
            state['step'] = torch.tensor(0.)
            state['exp_avg'] = torch.zeros_like(p, memory_format=torch.preserve_format)
            state['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.preserve_format)
        state['step'] += 1
        step = state['step'].item()  # CPU tensor: no device sync
        bias_correction1 = 1 - beta1 ** step
        bias_correction2 = 1 - beta2 ** step
        return state, group['lr'] / bias_correction1, math.sqrt(bias_correction2), beta1, beta2

    @torch.no_grad()
    def step_sparse(self, sparse_grads, gate="opacity"):
        """One step from COMPACT gradients (``diff_gaussian_rasterization.SparseGradients``, filled by the backward
        of ``render_post(..., sparse_grads=...)``): the rows the render touched are updated, every other row keeps
        its parameter and (stale) moments -- what ``step(relevant)`` does with
        ``relevant = (opacity.grad != 0).nonzero()`` (train_single.py:225-233), here without N-row gradient tensors
        and without a host round trip: the row list and its length stay on the device.  ``gate``: name of the
        compact gradient whose zero rows are skipped (the opacity gradient, as in the reference); None updates
        every touched row."""
        sg = sparse_grads
        if sg is None or not sg.ready:
            raise RuntimeError("step_sparse: the SparseGradients object has not been filled by a backward pass")
        by_ptr = {sg.params[k].data_ptr(): k for k in sg.grads if k in sg.params and sg.params[k].numel()}
        batch = []
        for group in self.param_groups:
            if group['amsgrad'] or group['maximize'] or group['weight_decay'] != 0 or group['capturable'] \
                    or group['foreach']:
                raise NotImplementedError("the fused sparse Adam implements the options the reference training uses")
            for p in group['params']:
                name = by_ptr.get(p.data_ptr())
                if name is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("fused sparse Adam needs contiguous float32 CUDA parameters (no CPU fallback)")
                state, step_size, bc2_sqrt, beta1, beta2 = self._update_terms(group, p)
                g = sg.grads[name]
                width = p.numel() // max(p.size(0), 1)
                batch.append(nat.AdamTensor(p.data_ptr(), g.data_ptr(), state['exp_avg'].data_ptr(),
                                            state['exp_avg_sq'].data_ptr(), width, step_size, bc2_sqrt, beta1, beta2,
                                            group['eps'], 1 - beta1, 1 - beta2))
        if not batch:
            return
        gate_t = sg.grads.get(gate) if gate is not None else None
        L = nat.lib()
        for i in range(0, len(batch), nat.ADAM_MAX_TENSORS):
            chunk = batch[i:i + nat.ADAM_MAX_TENSORS]
            arr = (nat.AdamTensor * len(chunk))(*chunk)
            nat.check(L.h3dgs_sparse_adam_compact(arr, len(chunk), sg.rows.data_ptr(), sg.count.data_ptr(),
                                                  sg.capacity, nat.ptr(gate_t), nat.stream_handle()),
                      "h3dgs_sparse_adam_compact")

    @torch.no_grad()
    def step_union(self, params, union_rows, union_count, union_packed, scale, row_map, capacity, out=None):
        """One step from the summed gradients of a data-parallel exchange (h3dgs_parallel.CompactRowExchange):
        ``union_packed`` [capacity, sum of widths] holds, for each of the ``union_count[0]`` (device) rows listed in
        ``union_rows``, the concatenated gradients of ``params`` in that order; ``scale`` (1 / world) turns the sum into
        the average.  ``row_map`` is the exchange's row -> slot scratch, restored to -1 by the kernel.  ``out``
        (sharded optimizer): device addresses (rows, packed, count) of the block that receives the updated parameter
        rows for the peers."""
        order = {p.data_ptr(): i for i, p in enumerate(params)}
        batch = [None] * len(params)
        for group in self.param_groups:
            if group['amsgrad'] or group['maximize'] or group['weight_decay'] != 0 or group['capturable'] \
                    or group['foreach']:
                raise NotImplementedError("the fused sparse Adam implements the options the reference training uses")
            for p in group['params']:
                i = order.get(p.data_ptr())
                if i is None:
                    continue
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("fused sparse Adam needs contiguous float32 CUDA parameters (no CPU fallback)")
                state, step_size, bc2_sqrt, beta1, beta2 = self._update_terms(group, p)
                width = p.numel() // max(p.size(0), 1)
                batch[i] = nat.AdamTensor(p.data_ptr(), None, state['exp_avg'].data_ptr(), state['exp_avg_sq'].data_ptr(),
                                          width, step_size, bc2_sqrt, beta1, beta2, group['eps'], 1 - beta1, 1 - beta2)
        if any(b is None for b in batch):
            raise RuntimeError("step_union: a tensor of the exchange is not a parameter of this optimizer")
        if len(batch) > nat.ADAM_MAX_TENSORS:
            raise NotImplementedError("step_union handles up to %d parameter tensors" % nat.ADAM_MAX_TENSORS)
        arr = (nat.AdamTensor * len(batch))(*batch)
        nat.check(nat.lib().h3dgs_dp_adam(arr, len(batch), union_rows.data_ptr(), union_count.data_ptr(),
                                          union_packed.data_ptr(), float(scale), nat.ptr(row_map), int(capacity),
                                          out[0] if out else None, out[1] if out else None, out[2] if out else None,
                                          nat.stream_handle()), "h3dgs_dp_adam")

    @torch.no_grad()
    def step(self, relevant, closure=None):
        """One optimisation step over the rows in ``relevant`` (1-D LongTensor of row indices; empty
        means every row)."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()

        batch = []  # (AdamTensor, rows) for one fused launch
        dev = None
        keep = []
        for group in self.param_groups:
            if group['amsgrad'] or group['maximize'] or group['weight_decay'] != 0 or group['capturable'] \
                    or group['foreach']:
                raise NotImplementedError(
                    "the fused sparse Adam implements the options the reference training uses "
                    "(amsgrad=False, maximize=False, weight_decay=0, capturable=False, foreach=None)")
            beta1, beta2 = group['betas']
            for p in group['params']:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError('Adam does not support sparse gradients, please consider SparseAdam instead')
                if not p.is_cuda or p.dtype != torch.float32:
                    raise RuntimeError("fused sparse Adam needs float32 CUDA parameters (no CPU fallback)")
                dev = p.device
                state = self.state[p]
                if len(state) == 0:
                    state['step'] = torch.tensor(0.)
                    state['exp_avg'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                state['step'] += 1
                step = state['step'].item()  # CPU tensor: no device sync
                bias_correction1 = 1 - beta1 ** step
                bias_correction2 = 1 - beta2 ** step
                step_size = group['lr'] / bias_correction1
                bias_correction2_sqrt = math.sqrt(bias_correction2)
                if not (p.is_contiguous() and state['exp_avg'].is_contiguous() and state['exp_avg_sq'].is_contiguous()):
                    raise RuntimeError("fused sparse Adam needs contiguous parameters and moments")
                grad = p.grad.contiguous()
                keep.append(grad)
                rows = p.size(0) if p.dim() > 0 else 1
                width = p.numel() // max(rows, 1)
                batch.append((nat.AdamTensor(p.data_ptr(), grad.data_ptr(), state['exp_avg'].data_ptr(),
                                             state['exp_avg_sq'].data_ptr(), width, step_size,
                                             bias_correction2_sqrt, beta1, beta2, group['eps'],
                                             1 - beta1, 1 - beta2), rows))
        if not batch:
            return loss

        L = nat.lib()
        if relevant is not None and relevant.numel() > 0:
            if relevant.dim() != 1:
                # train_coarse.py:133-134 passes the [K,2] result of a 2-D nonzero(): the reference's advanced
                # indexing (scene/OurAdam.py:290-337) then gathers every listed id (rows and the column zeros),
                # updates the copies and scatters them back, i.e. each DISTINCT id is updated exactly once
                relevant = relevant.reshape(-1).unique()
            # 1-D lists are used as given: sorted (nonzero() output) or duplicate-free; the kernel skips an id
            # equal to its predecessor, so sorted duplicates are applied once, as the reference's scatter does
            rel = relevant.to(device=dev, dtype=torch.int64).contiguous()
            rel_ptr, K = rel.data_ptr(), rel.numel()
        else:
            rel, rel_ptr, K = None, None, 0

        # tensors sharing a row count go out together, at most ADAM_MAX_TENSORS per launch
        by_rows = {}
        for t, rows in batch:
            by_rows.setdefault(rows, []).append(t)
        for rows, ts in by_rows.items():
            for i in range(0, len(ts), nat.ADAM_MAX_TENSORS):
                chunk = ts[i:i + nat.ADAM_MAX_TENSORS]
                arr = (nat.AdamTensor * len(chunk))(*chunk)
                nat.check(L.h3dgs_sparse_adam(arr, len(chunk), rel_ptr, K, rows, nat.stream_handle()),
                          "h3dgs_sparse_adam")
        return loss
