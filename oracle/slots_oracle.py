"""CPU oracle for the slot producers.  TEST INFRASTRUCTURE ONLY — imported by tests/ and nothing else.

numpy restatements of the reference tests' goldens, each pinned by tests/test_slot_ops.py against fixtures produced by the
reference's own functions (oracle/make_slots_golden.py):
  alloc_extend      tests/python/sgl_kernel_npu/test_alloc_extend_slot.py:9-85 (alloc_extend_pytorch) and :88-152 (native wrapper)
  cache_loc_assign  tests/python/sgl_kernel_npu/test_cache_assign.py:42-61; cache_loc_update is the reverse copy
                    (csrc/cache_location_assign/op_kernel/cache_loc_assign_kernel.cpp:161-174)
  assign_cache      tests/python/sgl_kernel_npu/test_inplace_assign_cache.py:27-46 with the op's own index tensors (:60-72)
"""
import numpy as np


def alloc_extend(pre_lens, seq_lens, last_loc, free_pages, page_size):
    """-> (out_indices int64 [sum(seq - pre)], pages taken).  Request i fills the rest of the page its prefix ends in from
    last_loc[i] + 1 on, then whole new pages and a last partial page from free_pages, consumed in request order."""
    pre, seq, last, free = (np.asarray(a, dtype=np.int64) for a in (pre_lens, seq_lens, last_loc, free_pages))
    out, page = [], 0
    for p0, s, ll in zip(pre, seq, last):
        first_new = -(-p0 // page_size)
        for p in range(p0, s):
            if p < first_new * page_size:
                out.append(ll + 1 + (p - p0))
            else:
                out.append(free[page + p // page_size - first_new] * page_size + p % page_size)
        if s > p0:
            page += -(-s // page_size) - first_new
    return np.asarray(out, dtype=np.int64), int(page)


def cache_loc_assign(req, token_pool, start, end, cache_loc):
    """token_pool[req[i], start_i:end_i] = cache_loc[packed run i]; returns the modified copy."""
    tp = np.array(token_pool, copy=True)
    off = 0
    for r, a, b in zip(np.asarray(req), np.asarray(start), np.asarray(end)):
        tp[r, a:b] = np.asarray(cache_loc)[off:off + (b - a)]
        off += b - a
    return tp


def cache_loc_update(req, token_pool, start, end, cache_loc):
    """cache_loc[packed run i] = token_pool[req[i], start_i:end_i]; returns the modified copy."""
    cl = np.array(cache_loc, copy=True)
    off = 0
    for r, a, b in zip(np.asarray(req), np.asarray(start), np.asarray(end)):
        cl[off:off + (b - a)] = np.asarray(token_pool)[r, a:b]
        off += b - a
    return cl


def assign_cache(out, src, dst_start, dst_end, src_start, src_end):
    """out[i, dst_start_i:dst_end_i] = src[src_start_i:src_end_i]; returns the modified copy."""
    o = np.array(out, copy=True)
    for i, (a, b, c, d) in enumerate(zip(*(np.asarray(x) for x in (dst_start, dst_end, src_start, src_end)))):
        n = min(b - a, d - c)
        o[i, a:a + n] = np.asarray(src)[c:c + n]
    return o
