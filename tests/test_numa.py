"""CPU: host placement helper of the host-buffer path (stage_indexorganized_b200/numa.py)."""
from stage_indexorganized_b200 import numa


def test_parse_cpulist():
    assert numa._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert numa._parse_cpulist("5") == {5}
    assert numa._parse_cpulist("") == set()
