#!/usr/bin/env python3
"""Transcribe the reference's own test tables into JSON fixtures.

Run in the build container, where /root/reference exists (it does not on the GPU box):

    python tests/golden/make_golden.py

Sources (the only tests in the reference tree):
  cmd/nvidia-device-plugin/replica_test.go:37-96   Test_prioritizeDevices  (15 cases)
  cmd/nvidia-device-plugin/replica_test.go:120-122 Test_stripReplicas      (3 cases)
  cmd/nvidia-device-plugin/nvidia_test.go:27-66    TestGetAdditionalXids   (10 cases)

The Go table literals are parsed textually (there is no Go toolchain here); the script
asserts the case counts so a parse slip cannot silently drop a case.
"""
import json
import os
import re
import sys

REF = os.environ.get("REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))


def go_strings(lit):
    """'[]string{"a", "b"}' or 'nil' -> list / None"""
    lit = lit.strip()
    if lit == "nil":
        return None
    m = re.fullmatch(r"\[\]string\{(.*)\}", lit, re.S)
    assert m, lit
    return re.findall(r'"((?:[^"\\]|\\.)*)"', m.group(1))


def parse_prioritize(src):
    body = src[src.index("func Test_prioritizeDevices"):src.index("func Test_stripReplicas")]
    pat = re.compile(
        r'\{"(?P<name>[^"]+)",[^\n]*\n\s*args\{(?P<avail>\[\]string\{[^}]*\}),\s*(?P<must>\[\]string\{[^}]*\}),\s*(?P<size>-?\d+)\},\s*\n'
        r'\s*(?P<want>\[\]string\{[^}]*\}|nil),\s*(?P<err>[^\n]+?),?\s*\n\s*\},')
    cases = []
    for m in pat.finditer(body):
        err = m.group("err").strip().rstrip(",")
        if err == "nil":
            e = {"kind": "none"}
        elif err.startswith("&NonUniqueError"):
            e = {"kind": "non_unique",
                 "message": "allocation resulted in non-unique devices due to requesting multiple GPU replicas and not having enough physical GPUs"}
        elif "no devices left to allocate" in err:
            e = {"kind": "no_devices_left", "message": "no devices left to allocate"}
        elif "mustIncludeDeviceIDs is missing" in err:
            arg = re.search(r',\s*"([^"]+)"\)$', err).group(1)
            e = {"kind": "must_missing", "arg": arg,
                 "message": "device '%s' in mustIncludeDeviceIDs is missing from availableDeviceIDs" % arg}
        else:
            raise AssertionError(err)
        cases.append({
            "name": m.group("name"),
            "available": go_strings(m.group("avail")),
            "must_include": go_strings(m.group("must")),
            "size": int(m.group("size")),
            "want": go_strings(m.group("want")),
            "error": e,
        })
    assert len(cases) == 15, len(cases)
    return cases


def parse_strip(src):
    body = src[src.index("func Test_stripReplicas"):]
    pat = re.compile(r'\{"(?P<name>[^"]+)",\s*args\{(?P<ids>\[\]string\{[^}]*\})\},\s*(?P<want>\[\]string\{[^}]*\})\},')
    cases = [{"name": m.group("name"), "ids": go_strings(m.group("ids")), "want": go_strings(m.group("want"))}
             for m in pat.finditer(body)]
    assert len(cases) == 3, len(cases)
    return cases


def parse_xids(src):
    body = src[src.index("testCases := []struct"):src.index("for i, tc := range testCases")]
    body = body[body.index("}{") + 2:]
    cases = []
    depth = 0
    cur = ""
    for ch in body:
        if ch == "{":
            if depth == 0:
                cur = ""
            else:
                cur += ch
            depth += 1
        elif ch == "}":
            depth -= 1
            if depth == 0:
                inp = re.search(r'input:\s*"((?:[^"\\]|\\.)*)"', cur)
                exp = re.search(r"expected:\s*\[\]uint64\{([^}]*)", cur)
                cases.append({
                    "input": inp.group(1) if inp else "",
                    "expected": [int(x) for x in exp.group(1).split(",") if x.strip()] if exp else [],
                })
            elif depth > 0:
                cur += ch
            else:
                break
        elif depth > 0:
            cur += ch
    assert len(cases) == 10, len(cases)
    return cases


def main():
    d = os.path.join(REF, "cmd", "nvidia-device-plugin")
    replica = open(os.path.join(d, "replica_test.go")).read()
    nvidia = open(os.path.join(d, "nvidia_test.go")).read()
    out = {
        "source": {
            "prioritize_devices": "cmd/nvidia-device-plugin/replica_test.go:37-96",
            "strip_replicas": "cmd/nvidia-device-plugin/replica_test.go:120-122",
            "additional_xids": "cmd/nvidia-device-plugin/nvidia_test.go:27-66",
        },
        "prioritize_devices": parse_prioritize(replica),
        "strip_replicas": parse_strip(replica),
        "additional_xids": parse_xids(nvidia),
    }
    path = os.path.join(HERE, "reference_tables.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1, sort_keys=False)
        f.write("\n")
    print("wrote", path, {k: len(v) for k, v in out.items() if isinstance(v, list)})


if __name__ == "__main__":
    sys.exit(main())
