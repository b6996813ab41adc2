"""Micro-benchmark of the bag-of-words transform kernel (csrc/bow.cu) on an ORBvoc-shaped tree
(k = 10, L = 6: 1 111 110 nodes, 10^6 words, 35.6 MB of node descriptors) and 512 frames x 2000 descriptors.
Prints one JSON line: descriptors/s device-resident (CUDA events on the vocabulary's stream) and through the host entry point."""
import ctypes as C
import importlib
import json
import os
import sys
import time
import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
bow = importlib.import_module("orb_slam2-chinese-annotation_b200.bow")
capi = importlib.import_module("orb_slam2-chinese-annotation_b200.capi")


def big_vocabulary(k=10, L=6, seed=0):
    """Breadth-first full tree (node order = level order, children of a node consecutive), vectorised."""
    rng = np.random.default_rng(seed)
    parent = [np.zeros(1, np.int32)]; desc = [np.zeros((1, 32), np.uint8)]; leaf = [np.zeros(1, np.uint8)]
    first = 0; count = 1
    for level in range(1, L + 1):
        pd = np.repeat(desc[-1], k, axis=0)
        mask = np.packbits(rng.random((len(pd), 256)) < (0.5 if level == 1 else 0.125), axis=1)
        d = (pd & ~mask) | (rng.integers(0, 256, pd.shape, dtype=np.uint8) & mask)
        parent.append(np.repeat(np.arange(first, first + count, dtype=np.int32), k)); desc.append(d)
        leaf.append(np.full(len(d), 1 if level == L else 0, np.uint8))
        first += count; count *= k
    n = first + count
    weight = np.zeros(n); weight[-count:] = rng.uniform(0.5, 12, count)
    return dict(k=k, L=L, scoring=0, weighting=0, parent=np.concatenate(parent), is_leaf=np.concatenate(leaf), desc=np.concatenate(desc), weight=weight)


def main():
    frames, per = 512, 2000
    n = frames * per
    voc = big_vocabulary()
    v = bow.ORBVocabulary(max_features=n)
    v.load_arrays(**voc)
    rng = np.random.default_rng(1)
    leaves = np.nonzero(voc["is_leaf"])[0]
    q = voc["desc"][rng.choice(leaves, n)] ^ np.packbits(rng.random((n, 256)) < 0.08, axis=1)
    L = capi.lib()
    st = torch.cuda.ExternalStream(L.orb_vocabulary_stream(v.h))
    dq = torch.from_numpy(q).cuda(); dw = torch.empty(n, dtype=torch.int32, device="cuda"); dn = torch.empty(n, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    def run():
        capi.check(L.orb_bow_transform_device(v.h, C.c_void_p(dq.data_ptr()), n, 4, C.c_void_p(dw.data_ptr()), C.c_void_p(dn.data_ptr())))
    for _ in range(3):
        run()
    capi.check(L.orb_vocabulary_sync(v.h))
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    K = 10
    with torch.cuda.stream(st):
        e0.record(st)
        for _ in range(K):
            run()
        e1.record(st)
    capi.check(L.orb_vocabulary_sync(v.h))
    ms = e0.elapsed_time(e1) / K
    t0 = time.perf_counter()
    for _ in range(3):
        w, nd = v.descend(q, 4)
    host_ms = (time.perf_counter() - t0) / 3 * 1e3
    assert np.array_equal(w, dw.cpu().numpy())
    # algorithmic work: L levels x k children x 8 popc per descriptor; bytes: L x k x 32 B gathered (L2-resident tree) + 32 B + 8 B
    print(json.dumps({"kernel": "k_bow_transform", "descriptors": n, "ms": ms, "descriptors_per_s": n / ms * 1e3, "frames_per_s_2000kp": frames / ms * 1e3,
                      "gathered_GBps": n * 6 * 10 * 32 / ms / 1e6, "popc_Gops": n * 6 * 10 * 8 / ms / 1e6, "host_entry_ms": host_ms,
                      "tree_MB": voc["desc"].nbytes / 1e6}))


if __name__ == "__main__":
    main()
