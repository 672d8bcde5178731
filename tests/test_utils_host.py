"""Host-side utilities that the bench line depends on, without a GPU: the clock verdict (utils/clocks.py) and the
CPU-locality binding report (utils/numa.py) with the NVML / sysfs sources replaced by fakes."""
import os

import pytest

from b200cnn.utils import clocks, numa


def test_clock_verdict():
    ok = {"sm_mhz": 1965, "sm_max_mhz": 1965, "reasons": []}
    assert not clocks.rejected(ok)
    assert not clocks.rejected({"sm_mhz": 1700, "sm_max_mhz": 1965, "reasons": ["sw_power_cap"]})     # kept and noted
    assert not clocks.rejected({"sm_mhz": 1300, "sm_max_mhz": 1965, "reasons": ["sw_power_cap"]})
    for bad in ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"):
        assert clocks.rejected({"sm_mhz": 1965, "sm_max_mhz": 1965, "reasons": [bad]})
    assert clocks.rejected({"sm_mhz": 1200, "sm_max_mhz": 1965, "reasons": []})                        # leftover lock
    assert not clocks.rejected({"sm_mhz": None, "sm_max_mhz": None, "reasons": []})                    # no sampler


def test_cpulist_parser():
    assert numa._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert numa._parse_cpulist("") == set()


@pytest.mark.skipif(not hasattr(os, "sched_setaffinity"), reason="needs sched_setaffinity")
def test_binding_report_always_says_why(monkeypatch):
    allowed = os.sched_getaffinity(0)
    calls = []
    monkeypatch.setattr(os, "sched_setaffinity", lambda pid, cpus: calls.append(set(cpus)))
    monkeypatch.delenv("B200_NO_NUMA_BIND", raising=False)
    # 1. every allowed CPU is local: nothing to narrow
    monkeypatch.setattr(numa, "_nvml_cpus", lambda idx: set(allowed))
    r = numa.bind_host_to_gpu("cuda:0")
    assert r["action"] == "none" and "nothing to narrow" in r["reason"] and r["source"] == "nvml" and not calls
    # 2. NVML fails, sysfs knows a strict subset: bound to it (when it is big enough to drive a GPU)
    def boom(idx):
        raise RuntimeError("nvml not available")
    monkeypatch.setattr(numa, "_nvml_cpus", boom)
    subset = set(sorted(allowed)[: max(len(allowed) // 2, 1)])
    monkeypatch.setattr(numa, "_sysfs_cpus", lambda idx: subset)
    r = numa.bind_host_to_gpu("cuda:0")
    if len(subset) >= 4 and subset != allowed:
        assert r["action"] == "bound" and r["source"] == "sysfs" and calls == [subset] and r["cpus_after"] == len(subset)
    else:
        assert r["action"] == "none" and r["reason"]
    # 3. neither source knows anything: the reason carries both errors
    monkeypatch.setattr(numa, "_sysfs_cpus", boom)
    r = numa.bind_host_to_gpu("cuda:0")
    assert r["action"] == "none" and "nvml: RuntimeError" in r["reason"] and "sysfs: RuntimeError" in r["reason"]
    # 4. switched off
    monkeypatch.setenv("B200_NO_NUMA_BIND", "1")
    assert "disabled" in numa.bind_host_to_gpu("cuda:0")["reason"]
