"""Test-only kernel table for the sharded module: the same five compute steps as
tdfo_b200.distributed.CudaKernelOps, answered by the CPU oracle on CPU tensors.  It lets the
world_size-2 gloo tests exercise the routing / collective logic without a GPU.  Lives under
tests/ on purpose: the product package has no CPU path."""
from types import SimpleNamespace

import numpy as np
import torch

import oracle as orc


def _np(t):
    return None if t is None else t.detach().cpu().numpy()


def _table(t):
    """Weight storage as the oracle takes it, sharing memory: float32, or uint16 bit patterns for bf16."""
    if t.dtype == torch.bfloat16:
        return t.detach().view(torch.int16).numpy().view(np.uint16)
    return t.detach().numpy()


class OracleKernelOps:
    def make_group(self, weights, feat_row_base, feat_rows, feat_out_col, out_stride, state0, state1, err_flag):
        return SimpleNamespace(weights=weights, feat_row_base=feat_row_base, feat_rows=feat_rows,
                               feat_out_col=feat_out_col, out_stride=out_stride, state0=state0, state1=state1,
                               err_flag=err_flag)

    def complete_cumsum(self, lengths):
        return torch.from_numpy(orc.complete_cumsum(_np(lengths)))

    def bucketize_row_wise(self, lengths, offsets, indices, block_sizes, F, B, world, err_flag):
        nl, no, ni, _, _, oob = orc.bucketize_row_wise(_np(offsets), _np(indices), None, _np(block_sizes), F, B,
                                                       world, want_permute=False)
        if oob:
            err_flag |= 1
        return torch.from_numpy(nl), torch.from_numpy(no), torch.from_numpy(ni)

    def bucketize_row_wise_with_permute(self, lengths, offsets, indices, block_sizes, F, B, world, err_flag):
        nl, no, ni, _, perm, oob = orc.bucketize_row_wise(_np(offsets), _np(indices), None, _np(block_sizes), F, B,
                                                          world, want_permute=True)
        if oob:
            err_flag |= 1
        return torch.from_numpy(nl), torch.from_numpy(no), torch.from_numpy(ni), torch.from_numpy(perm)

    def permute_2d(self, permute, B, lengths, offsets, indices, out_total):
        ol, oo, oi, _ = orc.permute_2d(_np(permute), B, _np(lengths), _np(offsets), _np(indices))
        return torch.from_numpy(ol), torch.from_numpy(oo), torch.from_numpy(oi)

    def embedding_forward(self, group, indices, offsets, B, pooling_mode, out, backward_scratch=None):
        F = group.feat_row_base.numel()
        w = np.ascontiguousarray(_table(group.weights))
        res, oob = orc.embedding_forward(w, w.shape[1], _np(indices), _np(offsets), None, _np(group.feat_row_base),
                                         _np(group.feat_rows), _np(group.feat_out_col), F, B, pooling_mode,
                                         out_stride=group.out_stride)
        out.copy_(torch.from_numpy(res))
        if oob:
            group.err_flag |= 1
        return out

    def embedding_backward_update(self, group, grad_out, indices, offsets, B, pooling_mode, opt, scratch=None,
                                  keys_in_scratch=False):
        F = group.feat_row_base.numel()
        w = _table(group.weights)                       # shares memory with the parameter
        o = orc.OptArgs(opt.optimizer, opt.learning_rate, opt.eps, opt.weight_decay, opt.beta1, opt.beta2, opt.step)
        orc.embedding_backward_update(w, w.shape[1], _np(group.state0) if group.state0 is None else group.state0.numpy(),
                                      None if group.state1 is None else group.state1.numpy(), _np(grad_out),
                                      _np(indices), _np(offsets), _np(group.feat_row_base), _np(group.feat_rows),
                                      _np(group.feat_out_col), F, B, pooling_mode, o)

    def copy_feature_blocks(self, src, dst, B, D, meta):
        m = _np(meta).reshape(4, -1)
        s, d = src.reshape(-1), dst.reshape(-1)
        for j in range(m.shape[1]):
            so, ss, do, ds = (int(x) for x in m[:, j])
            for b in range(B):
                d[do + b * ds: do + b * ds + D] = s[so + b * ss: so + b * ss + D]
