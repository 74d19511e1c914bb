"""libff_b200/numa.py: sysfs parsing and the no-op cases of the per-rank host placement."""
import os

from libff_b200 import numa


def test_cpulist_parsing():
    assert numa._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert numa._parse_cpulist("") == set()
    assert numa._parse_cpulist("5") == {5}


def test_unknown_device_node_is_none():
    assert numa.gpu_numa_node("ffff:ff:1f.0") is None
    assert numa.node_cpus(10 ** 6) == set()


def test_bind_without_a_gpu_changes_nothing():
    before = os.sched_getaffinity(0)
    info = numa.bind_to_gpu_node(0)
    assert info["bound"] is False and "why" in info
    assert os.sched_getaffinity(0) == before
