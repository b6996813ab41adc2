"""CPU tests: C-ABI library loads and exports every declared symbol, host-side helpers, oracle
self-consistency, and the world_size-2 gloo path of the frame sharding."""
import ctypes
import importlib
import os
import re
import subprocess
import sys
import numpy as np
import pytest
import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "orb_slam2-chinese-annotation_b200"
capi = importlib.import_module(PKG + ".capi")
synth = importlib.import_module(PKG + ".synth")
shard = importlib.import_module(PKG + ".shard")
views = importlib.import_module(PKG + ".views")


def test_library_exports_every_declared_symbol():
    if not os.path.exists(capi.LIB_PATH):
        capi.build()
    hdr = open(os.path.join(ROOT, "include", "orb_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(orb_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    lib = ctypes.CDLL(capi.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), "missing export " + name
    assert declared == set(capi.EXPORTS), declared ^ set(capi.EXPORTS)


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    ext = importlib.import_module(PKG + ".extractor")
    with pytest.raises(capi.OrbError) as e:
        ext.ORBextractor(1000, 1.2, 8, 20, 7)
    assert e.value.status == capi.ORB_ERR_CUDA
    mt = importlib.import_module(PKG + ".matcher")
    with pytest.raises(capi.OrbError):
        mt.ORBmatcher()
    bow = importlib.import_module(PKG + ".bow")
    with pytest.raises(capi.OrbError) as e:
        bow.ORBVocabulary().load_arrays(**synth.make_vocabulary(4, 2, 0))
    assert e.value.status == capi.ORB_ERR_CUDA
    rect = importlib.import_module(PKG + ".rectify")
    with pytest.raises(capi.OrbError) as e:
        rect.Rectifier(*synth.make_rectify_maps(64, 48, 0))
    assert e.value.status == capi.ORB_ERR_CUDA


def test_bow_assembly_matches_the_map_semantics():
    """Host assembly of BowVector / FeatureVector (bow.assemble) against a literal std::map-style restatement of
    TemplatedVocabulary::transform (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1146-1207) for all weighting / scoring kinds."""
    bow = importlib.import_module(PKG + ".bow")
    rng = np.random.default_rng(3)
    n, W = 500, 60
    word = rng.integers(0, W, n).astype(np.int32); node = (word // 7).astype(np.int32)
    ww = rng.uniform(0.1, 9.0, W); ww[rng.integers(0, W, 5)] = 0.0
    for weighting in range(4):
        for scoring in range(6):
            v, fv = {}, {}
            for i in range(n):
                w = ww[word[i]]
                if w > 0:
                    if weighting in (bow.TF_IDF, bow.TF):
                        v[int(word[i])] = v.get(int(word[i]), 0.0) + w
                    else:
                        v.setdefault(int(word[i]), w)
                    fv.setdefault(int(node[i]), []).append(i)
            keys = sorted(v)
            must = scoring != bow.DOT_PRODUCT
            if weighting in (bow.TF_IDF, bow.TF) and v and not must:
                for k in keys:
                    v[k] /= float(len(v))
            if must:
                norm = 0.0
                for k in keys:
                    norm += v[k] * v[k] if scoring == bow.L2_NORM else abs(v[k])
                if scoring == bow.L2_NORM:
                    norm = np.sqrt(norm)
                if norm > 0:
                    for k in keys:
                        v[k] /= norm
            ids, vals, f = bow.assemble(word, node, ww, weighting, scoring)
            assert list(ids) == keys and np.array_equal(vals, np.array([v[k] for k in keys])), (weighting, scoring)
            assert list(f.node_ids) == sorted(fv) and list(f.items) == [i for k in sorted(fv) for i in fv[k]]


def test_cpp_dropin_headers_compile_and_fail_loudly_without_device(tmp_path):
    """include/ORBextractor.h / include/ORBmatcher.h (the reference's class names over the C ABI) build into
    the drop-in driver; without a CUDA device the extractor throws instead of computing on the host."""
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "cpp")])
    exe = os.path.join(ROOT, "tests", "cpp", "build", "test_dropin")
    assert os.path.exists(exe)
    import torch
    if torch.cuda.is_available():
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import cppio
    fin = str(tmp_path / "e.in")
    cppio.write(fin, {"params": np.array([64, 64, 100, 8, 20, 7], np.int32), "scale": np.array([1.2], np.float32), "image": np.zeros(64 * 64, np.uint8)})
    r = subprocess.run([exe, "extract", fin, str(tmp_path / "e.out")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "no CUDA device" in r.stderr, (r.returncode, r.stderr)


def test_product_package_does_not_touch_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, PKG)):
        for f in files:
            if f.endswith((".py", ".cu", ".h", ".cuh", ".cpp", ".hpp")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liborb_oracle" not in src and "import orc" not in src and "oracle/" not in src, os.path.join(dirpath, f)


def test_synth_is_deterministic():
    a = synth.make_image(320, 200, 5); b = synth.make_image(320, 200, 5)
    assert np.array_equal(a, b) and a.dtype == np.uint8 and a.shape == (200, 320)
    assert not np.array_equal(a, synth.make_image(320, 200, 6))
    l, r = synth.make_stereo_pair(320, 200, 1)
    assert l.shape == r.shape == (200, 320)


def test_oracle_tables_match_survey():
    t = orc.Extractor(2000, 1.2, 8, 20, 7).tables()
    assert list(t["n_per_level"]) == [434, 362, 302, 251, 209, 175, 145, 122]
    assert list(t["umax"]) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert t["sf"][0] == 1.0 and abs(t["sf"][7] - 1.2 ** 7) < 1e-5
    t = orc.Extractor(1200, 1.2, 8, 20, 7).tables()
    assert list(t["n_per_level"]) == [261, 217, 181, 151, 126, 105, 87, 72]


def test_oracle_extract_basic_properties():
    img = synth.make_image(640, 480, 0)
    o = orc.Extractor(1000, 1.2, 8, 20, 7)
    kps, desc = o.extract(img)
    assert 1000 <= len(kps) <= 1000 + 3 * 8 and desc.shape == (len(kps), 32)
    assert np.all(np.diff(kps["octave"]) >= 0)                     # levels concatenated in order
    assert np.all((kps["angle"] >= 0) & (kps["angle"] < 360.0001))
    lv_sizes = [(640, 480), (533, 400), (444, 333), (370, 278), (309, 231), (257, 193), (214, 161), (179, 134)]
    for l, (w, h) in enumerate(lv_sizes):
        assert o.level(l).shape == (h, w)
        k = o.level_keypoints(l)
        assert np.all(k["x"] >= 19) and np.all(k["x"] < w - 19) and np.all(k["y"] >= 19) and np.all(k["y"] < h - 19)
    kps2, desc2 = o.extract(img)
    assert np.array_equal(kps, kps2) and np.array_equal(desc, desc2)   # deterministic (canonical quadtree tie-break)
    e, d = o.extract(np.empty((0, 0), np.uint8))
    assert len(e) == 0


def test_oracle_quadtree_properties():
    rng = np.random.default_rng(0)
    n = 3000
    pts = np.unique(np.stack([rng.integers(0, 1209, n), rng.integers(0, 344, n)], 1), axis=0)
    order = np.lexsort((pts[:, 0], pts[:, 1]))
    c = np.concatenate([pts[order], rng.integers(7, 200, (len(pts), 1))], 1).astype(np.int32)
    kept = orc.distribute(c, 16, 16 + 1209, 16, 16 + 344, 434)
    assert 434 <= len(kept) <= 434 + 3 and len(set(kept.tolist())) == len(kept)
    few = np.array([[10, 10, 30], [400, 20, 31], [700, 300, 32], [1000, 100, 33], [1200, 340, 34]], np.int32)
    assert sorted(orc.distribute(few, 16, 16 + 1209, 16, 16 + 344, 434).tolist()) == [0, 1, 2, 3, 4]   # stops without growth
    assert len(orc.distribute(c[:0], 16, 1225, 16, 360, 434)) == 0


def test_oracle_descriptor_distance_and_bruteforce():
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, (50, 32), dtype=np.uint8); b = rng.integers(0, 256, (40, 32), dtype=np.uint8)
    ref = np.unpackbits(a[:, None, :] ^ b[None, :, :], axis=2).sum(2)
    bi, bd, sd = orc.match_bruteforce(a, b)
    assert np.array_equal(bd, ref.min(1)) and np.array_equal(bi, ref.argmin(1))
    assert np.array_equal(sd, np.sort(ref, 1)[:, 1])
    assert orc.descriptor_distance(a[0], a[0]) == 0 and orc.descriptor_distance(a[0], ~a[0]) == 256


def test_oracle_grid_matches_numpy_statement():
    d = synth.make_frame_arrays(1500, 3)
    f = views.Frame(d["x"], d["y"], d["octave"], d["angle"], d["desc"], d["sf"], d["bounds"], d["u_right"])
    cs, ci = orc.assign_features_to_grid(f)
    px = np.floor((f.x - f.bounds[0]) * f.inv_w + np.float32(0.5)).astype(int)   # round half away, positive values
    py = np.floor((f.y - f.bounds[1]) * f.inv_h + np.float32(0.5)).astype(int)
    ok = (px >= 0) & (px < 64) & (py >= 0) & (py < 48)
    assert len(ci) == ok.sum()
    for c in (0, 100, 1500, 3071):
        exp = [i for i in range(f.N) if ok[i] and px[i] * 48 + py[i] == c]
        assert list(ci[cs[c]:cs[c + 1]]) == exp


def test_oracle_stereo_roundtrip_on_synthetic_disparity():
    w, h, fx = 640, 480, 500.0
    left, right = synth.make_stereo_pair(w, h, 7, fx=fx)
    ol = orc.Extractor(1000, 1.2, 8, 20, 7); orr = orc.Extractor(1000, 1.2, 8, 20, 7)
    kl, dl = ol.extract(left); kr, dr = orr.extract(right)
    ur, dp, sad, kept = orc.stereo_match(ol, orr, kl, dl, kr, dr, 40.0, 0.0, fx)
    m = ur >= 0
    assert kept == m.sum() and kept > 100
    disp = kl["x"][m] - ur[m]
    assert np.all(disp > 0) and np.all(disp < fx)
    assert np.allclose(dp[m], np.float32(40.0) / disp.astype(np.float32), rtol=1e-6)
    assert np.median(disp) > 2.0   # the synthetic disparity field is >= 2 px


def test_shard_range_partitions():
    for n in (0, 1, 7, 256, 1024):
        for world in (1, 2, 3, 8):
            spans = [shard.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.shard_range(4, 2, 2)


_WORKER = r'''
import importlib, os, sys
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
import numpy as np, torch.distributed as dist, orc
shard = importlib.import_module("orb_slam2-chinese-annotation_b200.shard")
synth = importlib.import_module("orb_slam2-chinese-annotation_b200.synth")
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
n = 5
lo, hi = shard.shard_range(n, rank, world)
o = orc.Extractor(300, 1.2, 8, 20, 7)   # stand-in for the per-GPU extractor: same sharding / gather path
local = []
for i in range(lo, hi):
    k, d = o.extract(synth.make_image(200, 120, i))
    local.append((i, len(k), int(d.astype(np.int64).sum())))
full = shard.gather_per_frame(local, n, dist)
assert [f[0] for f in full] == list(range(n)), full
if rank == 0:
    ref = []
    for i in range(n):
        k, d = o.extract(synth.make_image(200, 120, i))
        ref.append((i, len(k), int(d.astype(np.int64).sum())))
    assert full == ref
    print("GATHER_OK", len(full))
dist.barrier(); dist.destroy_process_group()
'''


def test_sharding_world_size_2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    port = 29500 + (os.getpid() % 2000)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(script), ROOT]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "GATHER_OK 5" in r.stdout


def test_multi_gpu_slices_cover_the_batch_and_fail_loudly_without_a_gpu():
    """orb_multi_slice (the library's frame sharding): contiguous blocks that tile [0, batch); orb_multi_create needs a device."""
    import importlib
    stereo = importlib.import_module("orb_slam2-chinese-annotation_b200.stereo")
    capi = importlib.import_module("orb_slam2-chinese-annotation_b200.capi")
    for batch in (0, 1, 7, 256, 1024, 1023):
        for parts in (1, 2, 3, 8):
            nxt = 0
            sizes = []
            for p in range(parts):
                f, c = stereo.multi_slice(batch, parts, p)
                assert f == nxt and c >= 0
                nxt += c; sizes.append(c)
            assert nxt == batch and max(sizes) - min(sizes) <= 1
    if capi.lib().orb_device_count() == 0:
        import pytest
        with pytest.raises(capi.OrbError):
            stereo.MultiGpuFrontEnd(1000, 1.2, 8, 20, 7, 640, 480, 4)
