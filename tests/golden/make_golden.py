"""Extracts the reference's own golden vectors for the coarse-graining hot path into a
compact fixture that can travel to the GPU box (where /root/reference does not exist).

Source: /root/reference/external/vcm/tests/_coarsen_restarts_regression_tests/reference/*.json
(7 regression tags of external/vcm/tests/test_coarsen_restarts.py:32-61; written by the
reference itself with `to_json`).  All four restart categories are kept
(fv_core.res, fv_tracer.res, fv_srf_wnd.res and sfc_data -- the "next" row N1 of SURVEY.md 8f).
Inputs are NOT stored: they regenerate bit-exactly from np.random.seed(0) (see
oracle/vcm_ref.regression_inputs, external/synth/synth/core.py:59-67).

Also stores the three known answers of external/vcm/tests/test_mappm.py:5-44.

Run here (container with /root/reference):  python tests/golden/make_golden.py
"""
import glob
import json
import os

import numpy as np

REF = "/root/reference/external/vcm/tests/_coarsen_restarts_regression_tests/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "coarsen_restarts_golden.npz")
CATEGORIES = ["fv_core.res", "fv_tracer.res", "fv_srf_wnd.res", "sfc_data"]


def main():
    arrays = {}
    for path in sorted(glob.glob(os.path.join(REF, "*.json"))):
        base = os.path.basename(path)[: -len(".json")]
        cat = next((c for c in CATEGORIES if base.endswith("-" + c)), None)
        if cat is None:
            continue
        tag = base[: -len(cat) - 1]
        d = json.load(open(path))
        for name, v in d["data_vars"].items():
            a = np.array(v["data"], dtype=np.float64)  # JSON null -> nan
            arrays[f"{tag}|{cat}|{name}"] = a
            arrays[f"{tag}|{cat}|{name}|dims"] = np.array(v["dims"])
    # external/vcm/tests/test_mappm.py known answers (inputs, expected, decimal)
    arrays["mappm_kat|interior|p_in"] = np.array([[0.0, 1.0, 2.0, 3.0, 4.0, 5.0]])
    arrays["mappm_kat|interior|f_in"] = np.array([[0.0, 1.0, 2.0, 3.0, 4.0]])
    arrays["mappm_kat|interior|p_out"] = np.array([[0.5, 1.2, 2.4, 2.8, 3.2, 4.5]])
    arrays["mappm_kat|interior|expected"] = np.array([[0.35, 1.3, 2.1, 2.5, 3.35]], dtype=np.float32)
    arrays["mappm_kat|oob|p_in"] = np.array([[1.0, 2.0, 3.0, 4.0, 5.0]])
    arrays["mappm_kat|oob|f_in"] = np.array([[1.5, 2.5, 3.5, 4.5]])
    arrays["mappm_kat|oob|p_out"] = np.array([[0.0, 2.5, 3.5, 4.5, 50.0]])
    arrays["mappm_kat|oob|expected"] = np.array([[1.5, 3.0, 4.0, 4.502747]], dtype=np.float32)
    arrays["mappm_kat|nan|p_in"] = np.array([[1.0, 2.0, 3.0, 2.0, 5.0]])
    arrays["mappm_kat|nan|f_in"] = np.array([[np.nan] * 4])
    arrays["mappm_kat|nan|p_out"] = np.array([[0.0, 2.5, 3.5, 4.5, 50.0]])
    arrays["mappm_kat|nan|expected"] = np.array([[np.nan] * 4], dtype=np.float32)
    np.savez_compressed(OUT, **arrays)
    print(f"wrote {OUT}: {len(arrays)} arrays, {os.path.getsize(OUT)} bytes")


if __name__ == "__main__":
    main()
