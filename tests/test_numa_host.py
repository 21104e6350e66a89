"""CPU: host placement helper (mam3slam_b200/numa.py) against a fake sysfs tree."""
import os

from mam3slam_b200 import numa


def test_parse_cpulist():
    assert numa.parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert numa.parse_cpulist("") == set()
    assert numa.parse_cpulist("5") == {5}


def test_node_and_cpus_from_sysfs(tmp_path):
    pci, node = tmp_path / "pci", tmp_path / "node"
    (pci / "0000:1b:00.0").mkdir(parents=True)
    (pci / "0000:1b:00.0" / "numa_node").write_text("1\n")
    (pci / "0000:9a:00.0").mkdir()
    (pci / "0000:9a:00.0" / "numa_node").write_text("-1\n")          # a host that does not say
    (node / "node1").mkdir(parents=True)
    (node / "node1" / "cpulist").write_text("48-95,144-191\n")
    assert numa.node_of_pci_device("0000:1B:00.0", str(pci)) == 1     # CUDA reports upper-case hex, sysfs is lower case
    assert numa.node_of_pci_device("0000:9a:00.0", str(pci)) is None
    assert numa.node_of_pci_device("0000:ff:00.0", str(pci)) is None
    cpus = numa.cpus_of_node(1, str(node))
    assert len(cpus) == 96 and 48 in cpus and 191 in cpus and 0 not in cpus
    assert numa.cpus_of_node(7, str(node)) == set()


def test_bind_without_a_device_leaves_the_mask_alone():
    before = os.sched_getaffinity(0)
    r = numa.bind_to_device_node(0)
    if not r["bound"]:
        assert os.sched_getaffinity(0) == before and r["reason"]
    else:                                                             # a GPU box: the mask shrank to one node's CPUs
        assert os.sched_getaffinity(0) <= before
        os.sched_setaffinity(0, before)
