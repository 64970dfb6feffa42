"""CPU-side checks of the C ABI: the library loads, exports every symbol the header
declares, fails loudly without a GPU, and its host-side string functions are bit-exact
with the reference's test tables and with the oracle."""
import json
import os
import re
import subprocess

import numpy as np
import pytest

from k8s_gpu_sharing_plugin_b200 import _lib as gpa

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_tables.json")))


@pytest.fixture(scope="module", autouse=True)
def built():
    import __graft_entry__ as entry
    entry.build_native()


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "gpualloc.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(gpa_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    syms = declared_symbols()
    assert len(syms) >= 15
    out = subprocess.check_output(["nm", "-D", "--defined-only", gpa.LIB_PATH]).decode()
    exported = set(re.findall(r" T (gpa_[a-z0-9_]+)", out))
    missing = [s for s in syms if s not in exported]
    assert not missing, missing
    L = gpa.lib()
    for s in syms:
        assert getattr(L, s) is not None


def test_library_is_built_for_sm_100a():
    out = subprocess.check_output(["cuobjdump", "-lelf", gpa.LIB_PATH]).decode()
    assert "sm_100a" in out, out


def test_no_cpu_fallback_create_fails_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(gpa.GpaError) as e:
        gpa.Context(0)
    assert e.value.code == gpa.ERR_CUDA


@pytest.mark.parametrize("case", GOLD["prioritize_devices"], ids=lambda c: c["name"])
def test_prioritize_devices_reference_table(case):
    got, kind, arg = gpa.prioritize_devices(case["available"], case["must_include"], case["size"])
    assert got == case["want"]
    assert kind == case["error"]["kind"]
    if kind == "must_missing":
        assert arg == case["error"]["arg"]


@pytest.mark.parametrize("case", GOLD["strip_replicas"], ids=lambda c: c["name"])
def test_strip_replicas_reference_table(case):
    assert gpa.strip_replicas(case["ids"]) == case["want"]


@pytest.mark.parametrize("case", GOLD["additional_xids"], ids=lambda c: repr(c["input"]))
def test_additional_xids_reference_table(case):
    assert gpa.get_additional_xids(case["input"]) == case["expected"]


def test_link_type_scores_match_oracle(oracle):
    for t in range(-2, 25):
        assert gpa.link_type_score(t) == oracle.link_type_score(t)


def _random_ids(rng, n_phys, replicas, p_free):
    phys = ["GPU-%08x-%d" % (int(rng.integers(1 << 31)), i) for i in range(n_phys)]
    ids = [p + "-replica-" + str(r) for p in phys for r in range(replicas) if rng.random() < p_free]
    rng.shuffle(ids)
    return [str(x) for x in ids]


@pytest.mark.parametrize("replicas", [1, 4, 16, 32])
def test_prioritize_devices_matches_oracle_on_fragmented_masks(oracle, replicas):
    """config 3 of SURVEY.md §8d in miniature: seeded fragmented availability, sizes 0..16,
    optional must-includes (valid, duplicated-physical, missing)."""
    rng = np.random.default_rng(300 + replicas)
    n = 0
    for trial in range(400):
        ids = _random_ids(rng, 8, replicas, float(rng.choice([0.1, 0.5, 0.9])))
        size = int(rng.choice([0, 1, 2, 3, 4, 8, 16]))
        must = []
        r = rng.random()
        if ids and r < 0.3:
            must = [str(x) for x in rng.choice(ids, size=min(len(ids), int(rng.integers(1, 4))), replace=False)]
        elif r < 0.35:
            must = ["GPU-missing-replica-0"]
        elif ids and r < 0.4:
            must = [ids[0].split("-replica-")[0] + "-replica-999"]
        want = oracle.prioritize_devices(ids, must, size)
        got = gpa.prioritize_devices(ids, must, size)
        assert got == want, (ids, must, size)
        # A9: the answer does not depend on the order kubelet lists the ids in
        perm = [str(x) for x in rng.permutation(ids)]
        assert gpa.prioritize_devices(perm, must, size) == want
        n += 1
    assert n == 400


def test_strip_replicas_matches_oracle_random(oracle):
    rng = np.random.default_rng(9)
    for _ in range(200):
        ids = _random_ids(rng, 8, 8, 0.5) + ["", "x", "a-replica-1-replica-2"][: int(rng.integers(0, 4))]
        assert gpa.strip_replicas(ids) == oracle.strip_replicas(ids)


def test_xids_match_oracle_random(oracle):
    rng = np.random.default_rng(3)
    alphabet = list("0123456789,,, -+x_") + [chr(9), "18446744073709551616", " ", ""]
    for _ in range(500):
        s = "".join(rng.choice(alphabet, size=int(rng.integers(0, 24))))
        assert gpa.get_additional_xids(s) == oracle.get_additional_xids(s), repr(s)


def test_prioritize_devices_with_unusual_replica_tails(oracle):
    """the groups are sorted by a packed 8-byte key of what follows "-replica-"; tails that do not fit (long, empty,
    no separator at all, a second separator) must fall back to the byte-wise string order of sort.Strings (replica.go:116)"""
    rng = np.random.default_rng(77)
    tails = ["-replica-0", "-replica-1", "-replica-10", "-replica-2", "-replica-", "-replica-123456789", "-replica-12345678",
             "-replica-a", "-replica-1-replica-2", "", "-replica-01", "-replica-1x", "-replica-\x7f"]
    gpus = ["GPU-%08d-aaaa-bbbb-cccc-dddddddddddd" % g for g in range(4)]
    for trial in range(400):
        ids = []
        for g in gpus:
            pick = rng.choice(len(tails), size=int(rng.integers(0, len(tails) + 1)), replace=False)
            ids += [g + tails[i] for i in pick]
        ids = [str(x) for x in rng.permutation(ids)]
        size = int(rng.choice([0, 1, 2, 3, 5, 8]))
        must = [str(x) for x in rng.choice(ids, size=min(len(ids), int(rng.integers(0, 3))), replace=False)] if ids else []
        assert gpa.prioritize_devices(ids, must, size) == oracle.prioritize_devices(ids, must, size), (ids, must, size)


def test_product_never_reaches_into_the_oracle():
    """oracle/ is test infrastructure: nothing in the package imports it, the library does not link it, and no product
    source names its symbols (ref_*)"""
    import ast
    pkg = os.path.join(ROOT, "k8s-gpu-sharing-plugin_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            tree = ast.parse(open(os.path.join(pkg, fn)).read())
            for node in ast.walk(tree):
                names = []
                if isinstance(node, ast.Import):
                    names = [a.name for a in node.names]
                elif isinstance(node, ast.ImportFrom):
                    names = [node.module or ""]
                assert not any(n.split(".")[0] == "oracle" for n in names), (fn, names)
    for fn in os.listdir(os.path.join(pkg, "csrc")):
        if not fn.endswith((".cu", ".cuh", ".cpp", ".h", ".inc")) and fn != "Makefile":
            continue
        src = open(os.path.join(pkg, "csrc", fn)).read()
        assert "liboracle" not in src and not re.search(r"\bref_[a-z_]+\s*\(", src), fn
    needed = subprocess.run(["readelf", "-d", os.path.join(pkg, "libgpualloc.so")], capture_output=True, text=True).stdout
    assert "NEEDED" in needed
    assert "liboracle" not in needed


def test_additional_xids_trim_every_unicode_space(oracle):
    """strings.TrimSpace (cmd/nvidia-device-plugin/nvidia.go:281) trims by unicode.IsSpace: not only ASCII white space and
    the Latin-1 NEL / NBSP but also U+1680, U+2000..U+200A, U+2028, U+2029, U+202F, U+205F and U+3000.  A space INSIDE a
    token is not trimmed and makes it malformed (skipped)."""
    s = " 13\u2003,\u300045\u00a0,\u168031\u2028\u2029,\u202f\u205f68\t\v,7\u20039,\u0085\n94\r\f, \u200b5"
    want = [13, 45, 31, 68, 94]     # "7<em space>9" and "<zero width space>5" (not White_Space) are malformed
    assert gpa.get_additional_xids(s) == want
    assert oracle.get_additional_xids(s) == want


def test_config3_fragmented_masks_match_the_oracle():
    """SURVEY.md 8d config 3 (tools/replica_bench.cpp, 300 of its 10 000 masks per case): 128 and 256 virtual devices,
    every replica free with p in {0.1, 0.5, 0.9}, sizes {1, 2, 3, 4, 8, 16}: prioritizeDevices natively against the
    oracle, ids, error kind and NonUniqueError alike"""
    import json
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.check_call(["make", "-C", os.path.join(root, "tools"), "-s", "bin/replica_bench"])
    out = subprocess.run([os.path.join(root, "tools", "bin", "replica_bench"), "300"], capture_output=True, text=True, timeout=300)
    r = json.loads(out.stdout)
    assert out.returncode == 0 and len(r["config3"]) == 6
    assert all(c["mismatches"] == 0 and c["decisions"] == 1800 for c in r["config3"])
    assert all(c["mismatches"] == 0 for c in r["cases"])
