"""Test-only stand-in for NVLink peer memory, so the peer-memory branch of the sharded module (the one
the GPUs run: `_peer_meta`, `_peer_grad_meta`, the push / landing-zone layout, the fused output
dist) can be exercised on CPU ranks at world sizes the 2-GPU tests do not reach.

"Peer memory" is a set of files under /dev/shm mapped by every rank; a "device pointer" is
(rank + 1) << 44 | buffer << 36 | byte offset, so the product code's pointer arithmetic (adding byte
offsets, the address / 4 form used for float destinations) works unchanged.  The five kernels that touch
peer memory are restated here in numpy from their contracts in include/tdfo_b200.h; everything else
comes from the oracle-backed table in oracle_ops.py.  Barriers are dist.barrier()."""
import os

import numpy as np
import torch
import torch.distributed as dist

import oracle as orc
from oracle_ops import OracleKernelOps, _np, _table


class FakeMem:
    def __init__(self, tag: str, rank: int, world: int):
        self.tag, self.rank, self.world = tag, rank, world
        self.count = 0
        self.maps = {}

    def _path(self, r, k):
        return f"/dev/shm/{self.tag}_{r}_{k}"

    def empty(self, n, dtype):
        """A zeroed tensor living in this rank's k-th shared file; returns (tensor, k)."""
        k = self.count
        self.count += 1
        np_dtype = {torch.int32: np.int32, torch.int64: np.int64, torch.float32: np.float32}[dtype]
        nbytes = max(int(n), 1) * np.dtype(np_dtype).itemsize
        with open(self._path(self.rank, k), "wb") as f:
            f.truncate(nbytes)
        mm = np.memmap(self._path(self.rank, k), dtype=np.uint8, mode="r+")
        self.maps[(self.rank, k)] = mm
        return torch.from_numpy(mm.view(np_dtype)), k

    def ptrs(self, k):
        return [((r + 1) << 44) | (k << 36) for r in range(self.world)]

    def view(self, addr: int, np_dtype, count: int):
        r, k, off = (addr >> 44) - 1, (addr >> 36) & 0xff, addr & ((1 << 36) - 1)
        assert 0 <= r < self.world, f"bad fake pointer {addr:#x}"
        mm = self.maps.get((r, k))
        if mm is None:
            mm = self.maps[(r, k)] = np.memmap(self._path(r, k), dtype=np.uint8, mode="r+")
        nb = count * np.dtype(np_dtype).itemsize
        assert off + nb <= mm.size, f"fake peer access out of bounds: buffer {k} of rank {r}, {off}+{nb} > {mm.size}"
        return mm[off:off + nb].view(np_dtype)

    def cleanup(self):
        self.maps.clear()
        for k in range(self.count):
            try:
                os.remove(self._path(self.rank, k))
            except OSError:
                pass


class _FakeEvent:
    def record(self, *a):
        pass

    def synchronize(self):
        pass


class FakePeerExchange:
    """Same buffers, in the same order, as tdfo_b200.distributed._PeerExchange."""

    def __init__(self, mem: FakeMem, pg, device, B, F_total, D, n_tw, n_rw, ids_capacity, n_cw=None, Dc=0, seq_cap=0):
        W = len(n_tw)
        n_cw = n_cw if n_cw is not None else [0] * W
        self.B, self.capacity, self.step, self._pg = B, int(ids_capacity), 0, pg
        self.seq_cap = int(seq_cap)
        self._mem = mem
        fl_max = max(a + b for a, b in zip(n_tw, n_cw)) + n_rw
        self._n_bags = fl_max * W * B
        z0 = self._make_zone()                              # zone 0 is the exchange object itself, as in the product
        for k_, v_ in vars(z0).items():
            setattr(self, k_, v_)
        self.zones = [self]
        self.out, self.out_ptrs = [], []
        for _ in range(2):
            t_, k = mem.empty(B * F_total * D, torch.float32)
            self.out.append(t_)
            self.out_ptrs.append(torch.tensor(mem.ptrs(k), dtype=torch.int64))
        self.grad, k = mem.empty(max(W * B * max(n_tw) * D, 4), torch.float32)
        self.grad_ptrs = mem.ptrs(k)
        self.grad_cw, self.grad_cw_ptrs = None, []
        if max(n_cw) > 0:
            self.grad_cw, k = mem.empty(max(W * B * max(n_cw) * Dc, 4), torch.float32)
            self.grad_cw_ptrs = mem.ptrs(k)
        self.rw_part, self.rw_part_ptrs, self.grad_rw, self.grad_rw_ptrs = None, [], None, []
        if n_rw > 0:
            self.rw_part, k = mem.empty(W * B * n_rw * D, torch.float32)
            self.rw_part_ptrs = mem.ptrs(k)
            self.grad_rw, k = mem.empty(W * B * n_rw * D, torch.float32)
            self.grad_rw_ptrs = mem.ptrs(k)
        self.seq_out, self.seq_out_ptrs, self.grad_seq, self.grad_seq_ptrs = [], [], None, []
        if self.seq_cap:
            for _ in range(2):
                t_, k = mem.empty(self.seq_cap * D, torch.float32)
                self.seq_out.append(t_)
                self.seq_out_ptrs.append(torch.tensor(mem.ptrs(k), dtype=torch.int64))
            self.grad_seq, k = mem.empty(self.capacity * D, torch.float32)
            self.grad_seq_ptrs = mem.ptrs(k)
        dist.barrier(group=pg)                              # every rank's files exist

    def _make_zone(self):
        from types import SimpleNamespace
        mem, z = self._mem, SimpleNamespace()
        z.len_x, k = mem.empty(max(self._n_bags, 4), torch.int32)
        z.len_ptrs = mem.ptrs(k)
        z.off_x, k = mem.empty(self._n_bags + 4, torch.int64)
        z.off_ptrs = mem.ptrs(k)
        z.ids_x, k = mem.empty(max(self.capacity, 4), torch.int64)
        z.ids_ptrs = mem.ptrs(k)
        z.counts_host = torch.zeros(3, dtype=torch.int64)
        z.counts_event = _FakeEvent()
        z.pos_x, z.pos_ptrs = None, []
        if self.seq_cap:
            z.pos_x, k = mem.empty(max(self.capacity, 4), torch.int64)
            z.pos_ptrs = mem.ptrs(k)
        return z

    def ensure_second_zone(self):
        if len(self.zones) < 2:
            self.zones.append(self._make_zone())
            dist.barrier(group=self._pg)                    # the new files exist everywhere

    def barrier(self, channel: int = 0):
        dist.barrier(group=self._pg)


class FakePeerOps(OracleKernelOps):
    def __init__(self, mem: FakeMem):
        self.mem = mem

    def make_peer_exchange(self, *args, **kwargs):
        return FakePeerExchange(self.mem, *args, **kwargs)

    # ---- tdfo_kjt_complete_cumsum into a caller buffer
    def complete_cumsum_into(self, lengths, out):
        res = orc.complete_cumsum(_np(lengths))
        out[: res.size].copy_(torch.from_numpy(res))
        return out

    # ---- tdfo_kjt_push_lengths: slot j copies B lengths from src_off[j] to the absolute address dst[j]
    def push_lengths(self, lengths, B, meta):
        m = _np(meta).reshape(2, -1)
        src = _np(lengths)
        for j in range(m.shape[1]):
            self.mem.view(int(m[1, j]), np.int32, B)[:] = src[int(m[0, j]): int(m[0, j]) + B]

    # ---- tdfo_kjt_push_ids: slot j streams one feature segment to where the owner's offsets say
    def push_ids(self, values, offsets, B, meta, capacity, err_flag):
        m = _np(meta).reshape(3, -1)
        vals, off = _np(values), _np(offsets)
        for j in range(m.shape[1]):
            seg = int(m[0, j])
            s0, s1 = int(off[seg * B]), int(off[(seg + 1) * B])
            if s1 <= s0:
                continue
            dst_off = int(self.mem.view(int(m[1, j]), np.int64, 1)[0])
            if dst_off < 0 or dst_off + (s1 - s0) > capacity:
                err_flag |= 4
                continue
            self.mem.view(int(m[2, j]) + 8 * dst_off, np.int64, s1 - s0)[:] = vals[s0:s1]

    # ---- tdfo_embedding_forward_peer: bag (f, gb) belongs to rank gb // peer_batch; its pooled row is
    # stored at out[rank] + ((gb % peer_batch) * out_stride + feat_out_col[f]) floats
    def embedding_forward_peer(self, group, indices, offsets, B, pooling_mode, peer_out, peer_batch,
                               backward_scratch=None, sort_pairs=None, expected_ids=0):
        F = group.feat_row_base.numel()
        w = np.ascontiguousarray(_table(group.weights))
        Dw = w.shape[1]
        local_col = (np.arange(F) * Dw).astype(np.int64)
        res, oob = orc.embedding_forward(w, Dw, _np(indices), _np(offsets), None, _np(group.feat_row_base),
                                         _np(group.feat_rows), local_col, F, B, pooling_mode, out_stride=F * Dw)
        if oob:
            group.err_flag |= 1
        ptrs, cols, stride = _np(peer_out), _np(group.feat_out_col), group.out_stride
        for s in range(B // peer_batch):
            for b in range(peer_batch):
                for f in range(F):
                    dst = self.mem.view(int(ptrs[s]) + 4 * (b * stride + int(cols[f])), np.float32, Dw)
                    dst[:] = res[s * peer_batch + b, f * Dw:(f + 1) * Dw]

    # ---- tdfo_copy_feature_blocks with dst == NULL: destinations are absolute addresses / 4
    def copy_feature_blocks(self, src, dst, B, D, meta):
        if dst is not None:
            return super().copy_feature_blocks(src, dst, B, D, meta)
        m = _np(meta).reshape(4, -1)
        s = _np(src).reshape(-1)
        for j in range(m.shape[1]):
            so, ss, do, ds = (int(x) for x in m[:, j])
            for b in range(B):
                self.mem.view(4 * (do + b * ds), np.float32, D)[:] = s[so + b * ss: so + b * ss + D]

    # ---- tdfo_reduce_partial_sums: out[b, out_col[k] + d] = sum over source slots, in rank order, fp32
    def reduce_partial_sums(self, parts, W, B, nslots, D, out_col, out, out_stride):
        p = _np(parts).reshape(W, B, nslots, D)
        acc = p[0].copy()
        for s in range(1, W):
            acc += p[s]
        o = out.view(B, out_stride)
        for k, c in enumerate(_np(out_col)):
            o[:, int(c): int(c) + D] = torch.from_numpy(acc[:, k])

    # ---- tdfo_kjt_scatter_positions: out[permute[i]] = positions_in[i] (or i), i < min(n, *n_dev)
    def scatter_positions(self, permute, positions_in, n, n_dev=None):
        if n_dev is not None:
            n = min(n, int(_np(n_dev)[0]))
        out = np.zeros(max(int(permute.numel()), 1), np.int64)
        src = np.arange(n, dtype=np.int64) if positions_in is None else _np(positions_in)[:n]
        out[_np(permute)[:n]] = src
        return torch.from_numpy(out)

    # ---- tdfo_kjt_push_rows: like push_ids, rows of D floats picked through `positions`
    def push_rows(self, rows, positions, offsets, B, meta, capacity, err_flag):
        m = _np(meta).reshape(3, -1)
        src, off = _np(rows), _np(offsets)
        D = src.shape[1]
        pos = None if positions is None else _np(positions)
        for j in range(m.shape[1]):
            seg = int(m[0, j])
            s0, s1 = int(off[seg * B]), int(off[(seg + 1) * B])
            if s1 <= s0:
                continue
            dst_off = int(self.mem.view(int(m[1, j]), np.int64, 1)[0])
            if dst_off < 0 or dst_off + (s1 - s0) > capacity:
                err_flag |= 4
                continue
            take = np.arange(s0, s1) if pos is None else pos[s0:s1]
            self.mem.view(int(m[2, j]) + 4 * dst_off * D, np.float32, (s1 - s0) * D)[:] = src[take].reshape(-1)

    # ---- tdfo_embedding_forward_seq_peer: landed id p of segment s -> row positions[p] of rank (s % world)'s buffer
    def embedding_forward_seq_peer(self, group, indices, offsets, B, positions, peer_out, world, sort_pairs=None):
        w = np.ascontiguousarray(_table(group.weights))
        if w.dtype == np.uint16:
            w = orc.bf16_to_f32(w)
        D = w.shape[1]
        idx, off, pos = _np(indices), _np(offsets), _np(positions)
        base, rows, ptrs = _np(group.feat_row_base), _np(group.feat_rows), _np(peer_out)
        S = base.size
        for s_ in range(S):
            for p in range(int(off[s_ * B]), int(off[(s_ + 1) * B])):
                i = int(idx[p])
                row = w[int(base[s_]) + i] if 0 <= i < int(rows[s_]) else np.zeros(D, np.float32)
                if not 0 <= i < int(rows[s_]):
                    group.err_flag |= 1
                self.mem.view(int(ptrs[s_ % world]) + 4 * int(pos[p]) * D, np.float32, D)[:] = row

    # ---- the id-only half of the backward has nothing to precompute for the oracle
    def backward_scratch_bytes(self, L, total_rows):
        return 256

    def embedding_backward_prepare(self, group, indices, offsets, B, pooling_mode, scratch, keys_in_scratch=False):
        pass
