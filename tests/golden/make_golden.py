"""Extracts the reference's own regression fixtures for the hot-path configs into small JSON files.

Run in the build container (reads /root/reference, which does not exist on the GPU box):
    python tests/golden/make_golden.py
Sources: tests/{2d,3d}_examples/<case>/regression_test_tool/*_Run_*_result.xml and
*_dtwdistance.xml.  Long series (3-D dam break: 2180 snapshots, twisting column: 3127) are kept in
full for the first run only and decimated for the others to keep the fixtures small.
"""
import json
import os
import re
import xml.etree.ElementTree as ET

REF = "/root/reference/tests"
OUT = os.path.dirname(os.path.abspath(__file__))

CASES = {
    "dambreak_2d": ("2d_examples/test_2d_dambreak", ["WaterBody_TotalMechanicalEnergy", "FluidObserver_Pressure"]),
    "dambreak_3d": ("3d_examples/test_3d_dambreak", ["WaterBody_TotalMechanicalEnergy", "FluidObserver_Pressure"]),
    "taylor_green_2d": ("2d_examples/test_2d_taylor_green", ["WaterBody_TotalKineticEnergy", "WaterBody_MaximumSpeed"]),
    "twisting_column_3d": ("3d_examples/test_3d_twisting_column", ["MyObserver_Position", "MyObserver_Velocity"]),
    # the CK / SYCL twins
    "dambreak_3d_ck": ("tests_sycl/3d_examples/test_3d_dambreak_sycl", ["WaterBody_TotalMechanicalEnergy", "FluidObserver_Pressure"]),
    "dambreak_2d_ck": ("tests_sycl/2d_examples/test_2d_dambreak_sycl", ["WaterBody_TotalMechanicalEnergy", "FluidObserver_Pressure"]),
}


def parse_value(s):
    s = s.strip()
    if s.startswith("~[") or s.startswith("["):
        return [float(x) for x in s.strip("~[]").split(",")]
    return float(s)


def parse_result(path):
    root = ET.parse(path).getroot()
    res = root.find("Result_Element")
    series = []
    for particle in res:
        items = sorted(((int(k.split("_")[1]), parse_value(v)) for k, v in particle.attrib.items() if k.startswith("snapshot_")))
        series.append([v for _, v in items])
    return series


def main():
    for case, (rel, quantities) in CASES.items():
        d = os.path.join(REF, rel, "regression_test_tool")
        out = {"source": f"tests/{rel}/regression_test_tool", "quantities": {}}
        for q in quantities:
            runs = sorted(f for f in os.listdir(d) if re.fullmatch(re.escape(q) + r"_Run_\d+_result\.xml", f))
            entry = {"runs": {}, "dtw_distance": None}
            for k, f in enumerate(runs):
                series = parse_result(os.path.join(d, f))
                n = len(series[0])
                stride = 1 if (k == 0 or n <= 100) else max(1, n // 100)
                entry["runs"][f] = {"n_snapshots": n, "stride": stride, "series": [s[::stride] for s in series]}
            dtw = os.path.join(d, q + "_dtwdistance.xml")
            if os.path.exists(dtw):
                node = ET.parse(dtw).getroot().find("DTWDistance")
                entry["dtw_distance"] = {k: parse_value(v) for k, v in node.attrib.items()}
            out["quantities"][q] = entry
        with open(os.path.join(OUT, case + ".json"), "w") as fh:
            json.dump(out, fh, separators=(",", ":"))
        print(case, {q: list(e["runs"]) for q, e in out["quantities"].items()}, os.path.getsize(os.path.join(OUT, case + ".json")))


if __name__ == "__main__":
    main()
