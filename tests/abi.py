"""Thin torch-tensor helpers over the raw C-ABI (include/mink_b200.h) for the
GPU parity tests.  Every call goes straight to libmink_b200.so via ctypes."""
import ctypes as C

import numpy as np
import torch

from mink_octtree_stablediffusion_b200 import _lib

DT = {torch.float32: 0, torch.float64: 1, torch.bfloat16: 2, torch.float16: 3}


def P(t):
    return C.c_void_p(t.data_ptr()) if t is not None and t.numel() > 0 else None


def S():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def HP(a):
    return a.ctypes.data_as(C.c_void_p)


def insert(coords_np):
    lib = _lib.load()
    c = torch.from_numpy(np.ascontiguousarray(coords_np)).cuda()
    n, ncols = c.shape
    cap = int(lib.mk_hash_capacity(n))
    table = torch.empty(cap * 16, dtype=torch.uint8, device="cuda")
    ws = torch.empty(int(lib.mk_insert_workspace_bytes(n)), dtype=torch.uint8, device="cuda")
    uniq = torch.full((max(n, 1),), -7, dtype=torch.int32, device="cuda")
    inv = torch.full((max(n, 1),), -7, dtype=torch.int32, device="cuda")
    cout = torch.empty((max(n, 1), ncols), dtype=torch.int32, device="cuda")
    meta = torch.zeros(2, dtype=torch.int64, device="cuda")
    _lib.check(lib.mk_map_insert(P(c), n, ncols, P(table), cap, P(uniq), P(inv), P(cout), P(ws),
                                 C.c_void_p(meta.data_ptr()), C.c_void_p(meta.data_ptr() + 8), S()))
    h = meta.cpu()
    nu, status = int(h[0]), int(h[1]) & 0xFFFFFFFF
    return dict(n_unique=nu, status=status, table=table, cap=cap, unique_index=uniq[:nu].cpu().numpy(),
                inverse=inv[:n].cpu().numpy(), coords=cout[:nu].cpu().numpy(), coords_dev=cout[:nu].contiguous())


def insert2(coords_np, unique=False, int64_maps=False):
    """mk_map_insert2 (fused dedup / compaction pass)."""
    lib = _lib.load()
    c = torch.from_numpy(np.ascontiguousarray(coords_np)).cuda()
    n, ncols = c.shape
    cap = int(lib.mk_hash_capacity(n))
    table = torch.empty(cap * 16, dtype=torch.uint8, device="cuda")
    ws = torch.empty(int(lib.mk_insert2_workspace_bytes(n)), dtype=torch.uint8, device="cuda")
    dt = torch.int64 if int64_maps else torch.int32
    uniq = torch.full((max(n, 1),), -7, dtype=dt, device="cuda")
    inv = torch.full((max(n, 1),), -7, dtype=dt, device="cuda")
    cout = torch.empty((max(n, 1), ncols), dtype=torch.int32, device="cuda")
    meta = torch.full((2,), -3, dtype=torch.int64, device="cuda")
    flags = (1 if unique else 0) | (2 if int64_maps else 0)
    _lib.check(lib.mk_map_insert2(P(c), n, ncols, P(table), cap, flags, P(uniq), P(inv), P(cout), P(ws),
                                  C.c_void_p(meta.data_ptr()), C.c_void_p(meta.data_ptr() + 8), S()))
    h = meta.cpu()
    nu, status = int(h[0]), int(h[1]) & 0xFFFFFFFF
    return dict(n_unique=nu, status=status, table=table, cap=cap, unique_index=uniq[:nu].cpu().numpy(),
                inverse=inv[:n].cpu().numpy(), coords=cout[:nu].cpu().numpy(), coords_dev=cout[:nu].contiguous())


def lookup(table, cap, query_np):
    lib = _lib.load()
    q = torch.from_numpy(np.ascontiguousarray(query_np)).cuda()
    rows = torch.empty(max(len(q), 1), dtype=torch.int32, device="cuda")
    _lib.check(lib.mk_map_lookup(P(table), cap, P(q), len(q), q.shape[1], P(rows), S()))
    return rows[:len(q)].cpu().numpy()


def kmap_build(out_coords_dev, in_table, in_cap, offsets_np):
    """-> (nbr [K, ld] int32 cuda, ld, counts np [K])"""
    lib = _lib.load()
    n_out, ncols = out_coords_dev.shape
    K = len(offsets_np)
    ld = max(128, (n_out + 127) // 128 * 128)
    nbr = torch.full((K, ld), -9, dtype=torch.int32, device="cuda")
    counts = torch.zeros(K, dtype=torch.int64, device="cuda")
    offs = np.ascontiguousarray(offsets_np, dtype=np.int32)
    _lib.check(lib.mk_kmap_build(P(out_coords_dev), n_out, ncols, P(in_table), in_cap, HP(offs), K, P(nbr), ld,
                                 C.c_void_p(counts.data_ptr()), S()))
    return nbr, ld, counts.cpu().numpy()


def kmap_pairs(nbr, n_out, ld, K):
    lib = _lib.load()
    ws = torch.empty(int(lib.mk_kmap_pairs_workspace_bytes(n_out, K)), dtype=torch.uint8, device="cuda")
    starts = torch.zeros(K + 1, dtype=torch.int64, device="cuda")
    _lib.check(lib.mk_kmap_pairs(P(nbr), n_out, ld, K, C.c_void_p(starts.data_ptr()), None, None, P(ws), S()))
    st = starts.cpu().numpy()
    Pn = int(st[-1])
    ir = torch.empty(max(Pn, 1), dtype=torch.int32, device="cuda")
    orr = torch.empty(max(Pn, 1), dtype=torch.int32, device="cuda")
    if Pn:
        _lib.check(lib.mk_kmap_pairs(P(nbr), n_out, ld, K, C.c_void_p(starts.data_ptr()), P(ir), P(orr), P(ws), S()))
    ir, orr = ir.cpu().numpy(), orr.cpu().numpy()
    return [(ir[st[k]:st[k + 1]].astype(np.int64), orr[st[k]:st[k + 1]].astype(np.int64)) for k in range(K)]


def kmap_invert(nbr, n_out, ld, K, n_in):
    lib = _lib.load()
    ld_in = max(128, (n_in + 127) // 128 * 128)
    inv = torch.empty((K, ld_in), dtype=torch.int32, device="cuda")
    _lib.check(lib.mk_kmap_invert(P(nbr), n_out, ld, K, P(inv), n_in, ld_in, S()))
    return inv, ld_in


def _ws(dtype, n_out, n_in, K, c_in, c_out):
    lib = _lib.load()
    nbytes = int(lib.mk_conv_workspace_bytes2(DT[dtype], n_out, n_in, K, c_in, c_out))
    return torch.empty(nbytes, dtype=torch.uint8, device="cuda"), nbytes


def conv_fwd(x, w, nbr, ld, n_out, transpose_w=0, kflip=0, c_out=None, mask=None, legacy=False):
    """mk_conv_fwd_ex (legacy=True: the size-less mk_conv_fwd, which never takes the fp32 split path)."""
    lib = _lib.load()
    K = w.shape[0]
    c_in = x.shape[1]
    if c_out is None:
        c_out = w.shape[1] if transpose_w else w.shape[2]
    y = torch.empty((n_out, c_out), dtype=x.dtype, device="cuda")
    ws, nbytes = _ws(x.dtype, n_out, x.shape[0], K, c_in, c_out)
    if legacy:
        _lib.check(lib.mk_conv_fwd(DT[x.dtype], P(x), P(w), transpose_w, kflip, P(nbr), P(mask), ld, n_out, x.shape[0], K,
                                   c_in, c_out, P(y), P(ws), S()))
    else:
        _lib.check(lib.mk_conv_fwd_ex(DT[x.dtype], P(x), P(w), transpose_w, kflip, P(nbr), P(mask), ld, n_out, x.shape[0],
                                      K, c_in, c_out, P(y), None, P(ws), nbytes, S()))
    return y


def tc_eligible(dtype, n_rows, K, c_in, c_out):
    return bool(_lib.load().mk_conv_tc_eligible(DT[dtype], n_rows, K, c_in, c_out))


def tile_order(nbr, n_rows, ld, K, coords_dev=None, block_log2=8):
    """(order [n_rows], table in tile order [K, ld])"""
    lib = _lib.load()
    order = torch.empty(max(n_rows, 1), dtype=torch.int32, device="cuda")
    out = torch.empty((K, ld), dtype=torch.int32, device="cuda")
    ws = torch.empty(int(lib.mk_kmap_tile_order_workspace_bytes(n_rows)), dtype=torch.uint8, device="cuda")
    _lib.check(lib.mk_kmap_tile_order(P(nbr), n_rows, ld, K, P(coords_dev), 0 if coords_dev is None else coords_dev.shape[1],
                                      block_log2, P(order), P(out), P(ws), S()))
    return order[:n_rows], out


def conv_fwd_rows(x, w, nbr, ld, n_out, out_rows, transpose_w=0, kflip=0, mask=None):
    lib = _lib.load()
    K = w.shape[0]
    c_in = x.shape[1]
    c_out = w.shape[1] if transpose_w else w.shape[2]
    y = torch.empty((n_out, c_out), dtype=x.dtype, device="cuda")
    ws, nbytes = _ws(x.dtype, n_out, x.shape[0], K, c_in, c_out)
    _lib.check(lib.mk_conv_fwd_ex(DT[x.dtype], P(x), P(w), transpose_w, kflip, P(nbr), P(mask), ld, n_out, x.shape[0],
                                  K, c_in, c_out, P(y), P(out_rows), P(ws), nbytes, S()))
    return y


def rows_gather(src, idx):
    lib = _lib.load()
    dst = torch.empty((idx.numel(), src.shape[1]), dtype=src.dtype, device="cuda")
    _lib.check(lib.mk_rows_gather(P(src), P(idx), idx.numel(), src.shape[1] * src.element_size(), P(dst), S()))
    return dst


def conv_wgrad(x, dy, nbr, ld, n_out, K, mask=None, dy_rows=None, dw_f32=False):
    lib = _lib.load()
    c_in, c_out = x.shape[1], dy.shape[1]
    dw = torch.empty((K, c_in, c_out), dtype=torch.float32 if dw_f32 else x.dtype, device="cuda")
    ws, nbytes = _ws(x.dtype, n_out, x.shape[0], K, c_in, c_out)
    _lib.check(lib.mk_conv_wgrad_ex(DT[x.dtype], P(x), P(dy), P(dy_rows), P(nbr), P(mask), ld, n_out, x.shape[0], K,
                                    c_in, c_out, P(dw), int(dw_f32), P(ws), nbytes, S()))
    return dw


def tile_mask(nbr, n_rows, ld, K):
    lib = _lib.load()
    W = int(lib.mk_kmap_tile_mask_words(K))
    ntiles = (n_rows + 127) // 128
    m = torch.empty((max(ntiles, 1), W), dtype=torch.int32, device="cuda")
    _lib.check(lib.mk_kmap_tile_mask(P(nbr), n_rows, ld, K, P(m), S()))
    return m[:ntiles]


def last_path():
    return int(_lib.load().mk_conv_last_path())


def force_simt(on):
    return int(_lib.load().mk_conv_force_simt(int(on)))
