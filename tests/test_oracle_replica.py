"""Pins the oracle's replica.go / nvidia.go restatements to the reference's own test tables
(tests/golden/reference_tables.json, transcribed by tests/golden/make_golden.py from
cmd/nvidia-device-plugin/replica_test.go and nvidia_test.go)."""
import json
import os

import pytest

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "reference_tables.json")))


@pytest.mark.parametrize("case", GOLD["prioritize_devices"], ids=lambda c: c["name"])
def test_prioritize_devices_reference_table(oracle, case):
    got, kind, arg = oracle.prioritize_devices(case["available"], case["must_include"], case["size"])
    assert got == case["want"]
    assert kind == case["error"]["kind"]
    if kind == "must_missing":
        assert arg == case["error"]["arg"]


@pytest.mark.parametrize("case", GOLD["strip_replicas"], ids=lambda c: c["name"])
def test_strip_replicas_reference_table(oracle, case):
    assert oracle.strip_replicas(case["ids"]) == case["want"]


@pytest.mark.parametrize("case", GOLD["additional_xids"], ids=lambda c: repr(c["input"]))
def test_additional_xids_reference_table(oracle, case):
    assert oracle.get_additional_xids(case["input"]) == case["expected"]


def test_prioritize_swap_remove_perturbs_group_order(oracle):
    # SURVEY.md §8c vector (3): remove() is swap-with-last (replica.go:56-60), so taking a-0
    # as a must-include leaves the group as [a-2, a-1] and the next pick is a-2.
    got, kind, _ = oracle.prioritize_devices(["a-replica-0", "a-replica-1", "a-replica-2"], ["a-replica-0"], 2)
    assert got == ["a-replica-0", "a-replica-2"] and kind == "non_unique"


def test_prioritize_panics_when_must_exceeds_size(oracle):
    # make([]string, len(must), size) with len > cap (replica.go:121)
    assert oracle.prioritize_devices(["a-replica-0", "a-replica-1"], ["a-replica-0", "a-replica-1"], 1)[1] == "panic"
    assert oracle.prioritize_devices(["a-replica-0"], [], -1)[1] == "panic"


def test_prioritize_result_sorted_bytewise(oracle):
    ids = ["g-replica-%d" % i for i in range(12)]
    got, kind, _ = oracle.prioritize_devices(ids, [], 3)
    # byte-wise: "...-10" < "...-11" < "...-2"; group[0] after sort is "-0", then "-1", "-10"
    assert got == ["g-replica-0", "g-replica-1", "g-replica-10"] and kind == "non_unique"


def test_strip_replica_splits_at_first_separator(oracle):
    assert oracle.strip_replicas(["x-replica-1-replica-2", "plain", ""]) == ["", "plain", "x"]


def test_xids_overflow_and_spaces(oracle):
    assert oracle.get_additional_xids("18446744073709551615, 18446744073709551616,\t7\n,+5,1_0") == [18446744073709551615, 7]
