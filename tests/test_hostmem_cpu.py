"""hostmem.py: NUMA placement helper of the host-buffer legs (CPU-only checks: parsing, no-op behaviour, affinity restored)."""
import os

from starkpack_winterfell_b200 import hostmem


def test_parse_cpulist():
    assert hostmem._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert hostmem._parse_cpulist("") == set()
    assert hostmem._parse_cpulist("5") == {5}


def test_bind_is_a_noop_without_a_device_and_restores_affinity():
    before = os.sched_getaffinity(0)
    with hostmem.bind_to_gpu_node(0) as b:
        inside = os.sched_getaffinity(0)
        assert inside <= before and inside
        assert isinstance(b.info, dict) and "bound" in b.info
    assert os.sched_getaffinity(0) == before


def test_node_cpus_of_missing_node():
    assert hostmem.node_cpus(10 ** 6) == set()
