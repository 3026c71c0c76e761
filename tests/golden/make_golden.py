"""Generate tests/golden/golden_v1.npz — known answers for the oracle and the CUDA path.

The reference (massquantity/tdfo) has no tests or vectors for this path and its
third-party kernels (torchrec / fbgemm_gpu) are not installable here, so the
answers come from the implementations those packages wrap or ship on CPU:

  * torch.nn.EmbeddingBag / nn.Embedding forward  (what an unsharded torchrec
    EmbeddingBagCollection / EmbeddingCollection calls per table),
  * torch autograd dense gradient + torch.optim.SGD / AdamW,
  * FBGEMM's `fbgemm::rowwise_sparse_adagrad_ref<long>` exported by libtorch_cpu.so,
  * hand-computed routing cases (bucketise, permute, jagged<->dense).

Run from the repo root:  python tests/golden/make_golden.py
It needs only torch + numpy (no GPU, no /root/reference at run time).
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path

import numpy as np
import torch

OUT = Path(__file__).resolve().parent / "golden_v1.npz"
G: dict[str, np.ndarray] = {}


def put(prefix: str, **arrays):
    for k, v in arrays.items():
        G[f"{prefix}.{k}"] = np.asarray(v)


def jagged_case(rng, F, B, rows, max_len, empty_every=4):
    lengths = rng.integers(0, max_len + 1, size=F * B).astype(np.int32)
    lengths[::empty_every] = 0                       # empty bags
    offsets = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64)
    idx = np.zeros(offsets[-1], dtype=np.int64)
    for f in range(F):
        s, e = offsets[f * B], offsets[(f + 1) * B]
        idx[s:e] = rng.integers(0, rows[f], size=e - s)
    if idx.size > 3:
        idx[1] = idx[0]                              # a repeated id inside the first non-empty bag
    return lengths, offsets, idx


def torch_ebc_forward(tables, idx, offsets, psw, F, B, mode):
    """tables: list of torch [rows_f, D] (requires_grad ok). Returns [B, F*D]."""
    outs = []
    for f in range(F):
        s, e = int(offsets[f * B]), int(offsets[(f + 1) * B])
        off = torch.from_numpy(offsets[f * B:(f + 1) * B + 1] - offsets[f * B])
        w = None if psw is None else torch.from_numpy(psw[s:e])
        outs.append(torch.nn.functional.embedding_bag(
            torch.from_numpy(idx[s:e]), tables[f], off, mode=mode, per_sample_weights=w,
            include_last_offset=True))
    return torch.cat(outs, dim=1)


def gen_forward(rng):
    F, B, D = 3, 6, 8
    rows = [11, 5, 7]
    lengths, offsets, idx = jagged_case(rng, F, B, rows, 5)
    tables = [torch.from_numpy(rng.uniform(-1, 1, size=(r, D)).astype(np.float32)) for r in rows]
    psw = rng.uniform(0.5, 1.5, size=idx.size).astype(np.float32)
    put("fwd", F=F, B=B, D=D, rows=rows, lengths=lengths, offsets=offsets, indices=idx, psw=psw,
        weights=torch.cat(tables).numpy(),
        out_sum=torch_ebc_forward(tables, idx, offsets, None, F, B, "sum").numpy(),
        out_mean=torch_ebc_forward(tables, idx, offsets, None, F, B, "mean").numpy(),
        out_wsum=torch_ebc_forward(tables, idx, offsets, psw, F, B, "sum").numpy())
    # sequence (EmbeddingCollection): nn.Embedding per table
    seq = torch.cat([torch.nn.functional.embedding(
        torch.from_numpy(idx[offsets[f * B]:offsets[(f + 1) * B]]), tables[f]) for f in range(F)])
    put("fwd", out_seq=seq.numpy())


def dense_grad(tables, idx, offsets, F, B, mode, grad_out):
    ts = [t.clone().requires_grad_(True) for t in tables]
    out = torch_ebc_forward(ts, idx, offsets, None, F, B, mode)
    (out * torch.from_numpy(grad_out)).sum().backward()
    return [t.grad for t in ts]


def gen_sgd(rng):
    F, B, D = 2, 8, 4
    rows = [6, 9]
    lengths, offsets, idx = jagged_case(rng, F, B, rows, 4)
    tables = [torch.from_numpy(rng.uniform(-1, 1, size=(r, D)).astype(np.float32)) for r in rows]
    grad_out = rng.normal(size=(B, F * D)).astype(np.float32)
    put("sgd", F=F, B=B, D=D, rows=rows, lengths=lengths, offsets=offsets, indices=idx,
        weights=torch.cat(tables).numpy(), grad_out=grad_out, lr=0.05)
    for mode in ("sum", "mean"):
        grads = dense_grad(tables, idx, offsets, F, B, mode, grad_out)
        new = []
        for t, g in zip(tables, grads):
            p = t.clone().requires_grad_(True)
            opt = torch.optim.SGD([p], lr=0.05)
            p.grad = g
            opt.step()
            new.append(p.detach())
        put("sgd", **{f"weights_after_{mode}": torch.cat(new).numpy()})


def fbgemm_rowwise_adagrad_ref():
    lib = ctypes.CDLL(os.path.join(os.path.dirname(torch.__file__), "lib", "libtorch_cpu.so"))
    fn = getattr(lib, "_ZN6fbgemm26rowwise_sparse_adagrad_refIlEEiiimPfPKfS1_PKT_fffPKdl")
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_void_p,
                   ctypes.c_void_p, ctypes.c_void_p, ctypes.c_float, ctypes.c_float, ctypes.c_float,
                   ctypes.c_void_p, ctypes.c_int64]
    return fn


def gen_rowwise_adagrad(rng):
    F, B, D = 2, 8, 8
    rows = [6, 9]
    total = sum(rows)
    lengths, offsets, idx = jagged_case(rng, F, B, rows, 4)
    tables = [torch.from_numpy(rng.uniform(-1, 1, size=(r, D)).astype(np.float32)) for r in rows]
    grad_out = rng.normal(size=(B, F * D)).astype(np.float32)
    h0 = rng.uniform(0.0, 0.5, size=total).astype(np.float32)
    lr, eps = 0.1, 1e-8
    grads = torch.cat(dense_grad(tables, idx, offsets, F, B, "sum", grad_out)).numpy()
    base = np.concatenate([[0], np.cumsum(rows)])
    touched = np.unique(np.concatenate(
        [idx[offsets[f * B]:offsets[(f + 1) * B]] + base[f] for f in range(F)])).astype(np.int64)
    fn = fbgemm_rowwise_adagrad_ref()
    put("rwa", F=F, B=B, D=D, rows=rows, lengths=lengths, offsets=offsets, indices=idx,
        weights=torch.cat(tables).numpy(), grad_out=grad_out, h0=h0, lr=lr, eps=eps)
    for wd in (0.0, 0.01):
        w = torch.cat(tables).numpy().copy()
        h = h0.copy()
        g = np.ascontiguousarray(grads[touched])
        # Caffe2 sign convention: w += step * g, so pass lr negated.
        rc = fn(touched.size, D, w.size, w.ctypes.data, g.ctypes.data, h.ctypes.data,
                touched.ctypes.data, eps, -lr, wd, None, 0)
        assert rc == touched.size, rc
        tag = "wd" if wd else "nowd"
        put("rwa", **{f"weights_after_{tag}": w, f"h_after_{tag}": h, f"wd_{tag}": wd})
    # one-row hand check of the formula itself (guards the sign / mean conventions above)
    w = np.array([[1.0, 2.0, 3.0, 4.0]], dtype=np.float32); h = np.array([0.25], dtype=np.float32)
    g = np.array([[0.5, -0.5, 1.0, 0.0]], dtype=np.float32); ix = np.array([0], dtype=np.int64)
    fn(1, 4, 4, w.ctypes.data, g.ctypes.data, h.ctypes.data, ix.ctypes.data, 0.0, -0.1, 0.0, None, 0)
    h_exp = 0.25 + (0.25 + 0.25 + 1.0) / 4
    np.testing.assert_allclose(h, [h_exp], rtol=1e-6)
    np.testing.assert_allclose(w, np.array([[1.0, 2.0, 3.0, 4.0]]) - 0.1 / np.sqrt(h_exp) * g, rtol=1e-6)


def gen_adam(rng):
    # every row is touched every step, so lazy (touched-rows-only) ADAM == dense AdamW
    F, B, D = 1, 12, 4
    rows = [4]
    steps = 3
    lr, eps, b1, b2 = 0.01, 1e-8, 0.9, 0.999
    w0 = rng.uniform(-1, 1, size=(rows[0], D)).astype(np.float32)
    put("adam", F=F, B=B, D=D, rows=rows, weights=w0, lr=lr, eps=eps, beta1=b1, beta2=b2, steps=steps)
    batches = []
    for s in range(steps):
        lengths = rng.integers(1, 4, size=F * B).astype(np.int32)
        offsets = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64)
        idx = rng.integers(0, rows[0], size=offsets[-1]).astype(np.int64)
        idx[:rows[0]] = np.arange(rows[0])
        grad_out = rng.normal(size=(B, F * D)).astype(np.float32)
        batches.append((lengths, offsets, idx, grad_out))
        put("adam", **{f"lengths{s}": lengths, f"offsets{s}": offsets, f"indices{s}": idx, f"grad_out{s}": grad_out})
    for wd in (0.0, 0.1):
        p = torch.from_numpy(w0.copy()).requires_grad_(True)
        opt = torch.optim.AdamW([p], lr=lr, betas=(b1, b2), eps=eps, weight_decay=wd)
        for s, (lengths, offsets, idx, grad_out) in enumerate(batches):
            opt.zero_grad()
            out = torch_ebc_forward([p], idx, offsets, None, F, B, "sum")
            (out * torch.from_numpy(grad_out)).sum().backward()
            opt.step()
            tag = "wd" if wd else "nowd"
            put("adam", **{f"weights_after_{tag}{s}": p.detach().numpy().copy(), f"wd_{tag}": wd})


def gen_routing():
    # SURVEY §8c KAT: W=2, rows=5 => block_size=3.  F=2 (second table rows=4 => block 2), B=2.
    #   bags (f,b): (0,0)=[4,0,3] (0,1)=[]  (1,0)=[1,3] (1,1)=[2,0,1]
    lengths = np.array([3, 0, 2, 3], dtype=np.int32)
    indices = np.array([4, 0, 3, 1, 3, 2, 0, 1], dtype=np.int64)
    weights = np.arange(8, dtype=np.float32) + 0.5
    block_sizes = np.array([3, 2], dtype=np.int64)
    # bucket 0: (0,0)=[0] (0,1)=[] (1,0)=[1] (1,1)=[0,1] ; bucket 1: (0,0)=[4-3,3-3]=[1,0] (0,1)=[] (1,0)=[3-2]=[1] (1,1)=[2-2]=[0]
    new_lengths = np.array([1, 0, 1, 2, 2, 0, 1, 1], dtype=np.int32)
    new_indices = np.array([0, 1, 0, 1, 1, 0, 1, 0], dtype=np.int64)
    # original position -> slot
    unbucketize = np.array([4, 0, 5, 1, 6, 7, 2, 3], dtype=np.int64)
    new_weights = np.zeros(8, dtype=np.float32)
    new_weights[unbucketize] = weights
    put("bucketize", F=2, B=2, world=2, lengths=lengths, indices=indices, weights=weights,
        block_sizes=block_sizes, new_lengths=new_lengths, new_indices=new_indices,
        new_weights=new_weights, unbucketize_permute=unbucketize)
    # permute_2d: 3 segments of B=2 bags, out = [seg2, seg0, seg2]
    in_lengths = np.array([1, 2, 0, 3, 2, 1], dtype=np.int32)
    in_indices = np.arange(10, 19, dtype=np.int64)      # seg0: [10 | 11 12], seg1: [ | 13 14 15], seg2: [16 17 | 18]
    put("permute", B=2, permute=np.array([2, 0, 2], dtype=np.int32), in_lengths=in_lengths,
        in_indices=in_indices, out_lengths=np.array([2, 1, 1, 2, 2, 1], dtype=np.int32),
        out_indices=np.array([16, 17, 18, 10, 11, 12, 16, 17, 18], dtype=np.int64))
    # jagged <-> dense: B=3, D=2, max_len=2; lengths [1,3,0]
    values = np.arange(8, dtype=np.float32).reshape(4, 2) + 1
    offsets = np.array([0, 1, 4, 4], dtype=np.int64)
    dense = np.zeros((3, 2, 2), dtype=np.float32)
    dense[0, 0] = values[0]; dense[1, 0] = values[1]; dense[1, 1] = values[2]
    back = values.copy(); back[3] = 0                    # row 3 (t=2 >= max_len) was truncated
    put("jagged", values=values, offsets=offsets, max_len=2, dense=dense, back=back,
        ids=np.array([7, 8, 9, 10], dtype=np.int64),
        padded=np.array([[7, -1, -1], [8, 9, 10], [-1, -1, -1]], dtype=np.int64), pad=-1, desired=3)


def main():
    torch.manual_seed(0)
    torch.set_num_threads(1)
    rng = np.random.default_rng(20261106)
    gen_forward(rng)
    gen_sgd(rng)
    gen_rowwise_adagrad(rng)
    gen_adam(rng)
    gen_routing()
    np.savez_compressed(OUT, **G)
    print(f"wrote {OUT} ({OUT.stat().st_size} bytes, {len(G)} arrays)")


if __name__ == "__main__":
    main()
