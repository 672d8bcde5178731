"""Warehouse ETL over the three CSV schemas + run logs, derived stats, speed-up, exports."""
import os
import subprocess
import sys

import pandas as pd
import pytest

from b200cnn.harness import log_analysis as la
from b200cnn.harness.suite import CSV_HEADER

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _make_tree(tmp):
    logs = tmp / "final_project" / "logs" / "s1"
    logs.mkdir(parents=True)
    rows = [("V1 Serial", 1, 800.0), ("V1 Serial", 1, 760.0), ("V2 2.2-scatter-halo", 2, 400.0),
            ("V2 2.2-scatter-halo", 4, 250.0), ("V4 MPI+CUDA", 1, 180.0), ("V4 MPI+CUDA", 4, 100.0)]
    df = pd.DataFrame([{**{h: "" for h in CSV_HEADER}, "EntryTimestamp": "2026-01-01T00:00:00",
                        "ProjectVariant": v, "NumProcesses": n, "ExecutionTime_ms": t} for v, n, t in rows])
    df.to_csv(logs / "summary_report_1.csv", index=False)
    pd.DataFrame({"Timestamp": ["2025-05-06 16:08:35"], "Version": ["V3 CUDA Only"], "NP": [1],
                  "Time_ms": [194.0]}).to_csv(tmp / "legacy.csv", index=False)
    pd.DataFrame({"timestamp": ["2025-05-06"], "version": ["V3 CUDA Only"], "np": [1],
                  "time_ms": [488.0]}).to_csv(tmp / "lower.csv", index=False)
    pd.DataFrame({"foo": [1]}).to_csv(tmp / "unknown.csv", index=False)
    (logs / "run_v4_mpi_cuda_np2.log").write_text("AlexNet MPI+CUDA Forward Pass completed in 150 ms\n"
                                                  "ExecutionTime_ms: 150.5\n")
    (logs / "run_v1_serial_np1.log").write_text("AlexNet Serial Forward Pass completed in 742 ms\n")
    (tmp / "k.cu").write_text("#include <x>\n__global__ void k(){}\nvoid host(int a) {\n}\n")
    return tmp


def test_ingest_stats_speedup_export(tmp_path):
    root = _make_tree(tmp_path)
    wh = la.Warehouse(tmp_path / "wh.sqlite")
    n = wh.ingest(root)
    assert n["summary_rows"] == 8 and n["run_logs"] == 1 and n["sources"] == 1
    assert wh.ingest(root)["files"] == 0                      # SHA-1 dedup
    st = wh.run_stats().set_index(["version", "np"])
    assert st.loc[("V1 Serial", 1), "n"] == 2
    assert abs(st.loc[("V1 Serial", 1), "mean_s"] - 0.78) < 1e-9
    assert abs(st.loc[("V3 CUDA Only", 1), "median_s"] - 0.341) < 1e-9
    sp = wh.speedup().set_index(["version", "np"])
    assert abs(sp.loc[("V2 2.2-scatter-halo", 4), "S"] - 0.76 / 0.25) < 1e-9
    assert abs(sp.loc[("V2 2.2-scatter-halo", 4), "E"] - 0.76 / 0.25 / 4) < 1e-9
    src = wh.df("SELECT * FROM source_stats")
    assert int(src.cuda_kernel_cnt[0]) == 1
    rl = wh.df("SELECT * FROM run_logs")
    assert rl.version[0] == "v4" and int(rl.np[0]) == 2 and abs(rl.total_time_s[0] - 0.1505) < 1e-9


def test_cli_roundtrip(tmp_path):
    root = _make_tree(tmp_path)
    db = str(tmp_path / "w.sqlite")
    cli = [sys.executable, os.path.join(ROOT, "log_analysis.py"), "--db", db]
    assert subprocess.run(cli + ["ingest", "--root", str(root), "--rebuild"], capture_output=True).returncode == 0
    out = subprocess.run(cli + ["speedup"], capture_output=True, text=True)
    assert out.returncode == 0 and "V4 MPI+CUDA" in out.stdout
    assert subprocess.run(cli + ["plot", "speedup", "efficiency", "--outdir", str(tmp_path / "plots")],
                          capture_output=True).returncode == 0
    assert (tmp_path / "plots" / "efficiency.svg").read_text().startswith("<svg")
    for ext in ("csv", "parquet"):
        of = tmp_path / f"best.{ext}"
        assert subprocess.run(cli + ["export", "best_runs", str(of)], capture_output=True).returncode == 0
        assert of.exists()
    assert subprocess.run(cli + ["export", "nope", str(tmp_path / "x.csv")], capture_output=True).returncode == 1


def test_canonical_names_and_time_ladder():
    assert la.canonical_version_name("V2 2.2-scatter-halo") == "V2.2 ScatterHalo"
    assert la.canonical_version_name("v2_2.1_broadcast_all") == "V2.1 BroadcastAll"
    assert la.canonical_version_name("V3 CUDA Only") == "V3 CUDA"
    assert la.canonical_version_name("V4 MPI+CUDA") == "V4 MPI+CUDA"
    assert la.canonical_version_name("V5 MPI+CUDA-Aware") == "V5 MPI+CUDA-Aware"
    assert la.time_from_log("Execution Time: 286.9 ms") == pytest.approx(0.2869)
    assert la.time_from_log("pid=0: Total PARALLEL time: 1.5 seconds.") == pytest.approx(1.5)


def test_locperf_and_legacy_graphs(tmp_path):
    root = _make_tree(tmp_path)
    (tmp_path / "final_project").mkdir(exist_ok=True)
    v3 = tmp_path / "final_project" / "v3_cuda_only"
    v3.mkdir()
    (v3 / "layers_cuda.cu").write_text("__global__ void a(){}\n__global__ void b(){}\n")
    db = str(tmp_path / "w2.sqlite")
    cli = [sys.executable, os.path.join(ROOT, "log_analysis.py"), "--db", db]
    assert subprocess.run(cli + ["ingest", "--root", str(root)], capture_output=True).returncode == 0
    out = subprocess.run(cli + ["locperf"], capture_output=True, text=True)
    assert out.returncode == 0 and "V3 CUDA" in out.stdout
    r = subprocess.run(cli + ["legacy-graphs", str(tmp_path / "legacy.csv"), "--outdir", str(tmp_path / "lg")],
                       capture_output=True, text=True)
    assert r.returncode == 0 and (tmp_path / "lg" / "legacy_times.svg").exists()


def test_v1_python_runner_equals_native_binary(tmp_path):
    """Same seed -> same FNV digest from the Python V1 runner and the native V1 `template`."""
    import re
    vdir = os.path.join(ROOT, "final_project", "v1_serial")
    assert subprocess.run(["make"], cwd=vdir, capture_output=True).returncode == 0
    nat = subprocess.run(["./template", "--seed", "77"], cwd=vdir, capture_output=True, text=True).stdout
    py = subprocess.run([sys.executable, "-m", "b200cnn.runners.v1", "--seed", "77"], cwd=ROOT, capture_output=True,
                        text=True, env=dict(os.environ, PYTHONPATH=ROOT)).stdout
    f = lambda t: re.search(r"Output FNV1a64: (0x[0-9a-f]+)", t).group(1)
    assert f(nat) == f(py)
    assert nat.splitlines()[0] == py.splitlines()[0] == "--- AlexNet Serial CPU Version (V1) ---"


def test_scorecard_pools_canonical_names(tmp_path):
    """Harness CSV labels and run-log tags of the same version land in one (version, np) group."""
    root = _make_tree(tmp_path)
    wh = la.Warehouse(tmp_path / "wh.sqlite")
    wh.ingest(root)
    raw = wh.run_stats().set_index(["version", "np"])
    assert ("v4", 2) in raw.index and ("V4 MPI+CUDA", 1) in raw.index          # separate labels
    st = wh.run_stats(canonical=True).set_index(["version", "np"])
    assert ("V4 MPI+CUDA", 2) in st.index and ("v4", 2) not in st.index
    assert st.loc[("V1 Serial", 1), "n"] == 2                                  # the two CSV rows (V1's log has no np != 1 row)
    sp = wh.speedup("median", canonical=True).set_index(["version", "np"])
    assert abs(sp.loc[("V2.2 ScatterHalo", 4), "S"] - 0.78 / 0.25) < 1e-9
    assert abs(sp.loc[("V4 MPI+CUDA", 2), "E"] - 0.78 / 0.1505 / 2) < 1e-9


def test_sql_views_and_notebook_figures(tmp_path):
    """The notebook's per-version `speedup` / `efficiency` SQL views (ref analysis.md:222-240) and its CV, radar
    and LOC-vs-performance figures (ref analysis.md:2305-2460), generated as SVG from the warehouse."""
    root = _make_tree(tmp_path)
    db = tmp_path / "wh.sqlite"
    wh = la.Warehouse(db)
    wh.ingest(root)
    sp = wh.df("SELECT * FROM speedup WHERE version = 'V4 MPI+CUDA' ORDER BY np")
    assert list(sp.np) == [1, 4] and abs(sp.S.iloc[0] - 1.0) < 1e-12          # per-version baseline: S(np=1) = 1
    assert abs(sp.S.iloc[1] - 1.8) < 1e-9                                     # 0.180 s / 0.100 s
    ef = wh.df("SELECT E FROM efficiency WHERE version = 'V4 MPI+CUDA' AND np = 4")
    assert abs(ef.E.iloc[0] - 0.45) < 1e-9
    rt = la.radar_table(wh)
    assert {"np1_perf_norm", "max_speedup_norm", "max_efficiency_norm", "log10_loc_norm"} <= set(rt.columns)
    out = tmp_path / "plots"
    rc = la.main(["--db", str(db), "plot", "cv", "radar", "locperf", "--outdir", str(out)])
    assert rc == 0
    for f in ("runtime_cv.svg", "radar.svg"):
        txt = (out / f).read_text()
        assert txt.startswith("<svg") and txt.rstrip().endswith("</svg>")
    assert "V1 Serial" in (out / "radar.svg").read_text()
    assert "n=2" in (out / "runtime_cv.svg").read_text()                       # V1 Serial np=1 has two runs


def test_analysis_notebook_cells_execute(tmp_path, monkeypatch):
    """analysis.ipynb (counterpart of the reference's notebook, generated by scripts/make_notebook.py) is valid
    nbformat-4 JSON, is what the generator writes, and every code cell runs in order on a small log tree."""
    import json
    import subprocess
    import sys
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    nb = json.load(open(os.path.join(repo, "analysis.ipynb")))
    assert nb["nbformat"] == 4 and [c["cell_type"] for c in nb["cells"]].count("code") >= 8
    fresh = tmp_path / "fresh.ipynb"
    subprocess.run([sys.executable, os.path.join(repo, "scripts", "make_notebook.py"), str(fresh)], check=True,
                   capture_output=True)
    assert json.load(open(fresh)) == nb, "analysis.ipynb is stale: run python scripts/make_notebook.py"
    root = _make_tree(tmp_path)
    monkeypatch.setenv("B200_ANALYSIS_ROOT", repo)
    monkeypatch.setenv("B200_ANALYSIS_LOGS", str(root))
    monkeypatch.setenv("B200_ANALYSIS_OUT", str(tmp_path / "nb_out"))
    env = {"__name__": "__notebook__"}
    for c in nb["cells"]:
        if c["cell_type"] == "code":
            exec(compile("".join(c["source"]), "<analysis.ipynb>", "exec"), env)
    out = tmp_path / "nb_out"
    for f in ("best_runs.csv", "best_runs.md", "stats.csv", "all_runs.csv", "stats.parquet", "radar_table.csv"):
        assert (out / f).stat().st_size > 0, f
    plots = {p.name for p in (out / "plots").glob("*.svg")}
    assert {"efficiency.svg", "runtime_cv.svg", "radar.svg"} <= plots and any(p.startswith("speedup_") for p in plots)
    assert not env["top"].empty and not env["radar"].empty and "S" in env["sp_best"].columns
