"""Log-line contract and harness parser round trips."""
import os
import re
import subprocess
import sys

import pytest

from b200cnn.harness import suite
from b200cnn.utils import logfmt as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cxx_float_matches_iostream():
    assert L.cxx_float(44.415199) == "44.4152"
    assert L.cxx_float(0.058493614) == "0.0584936"
    assert L.cxx_float(87.12) == "87.12"
    assert L.cxx_float(1e-5) == "1e-05"
    assert L.cxx_float(123456789.0) == "1.23457e+08"


def test_first_ms_wins_rule():
    # extras never use the "<num> ms" form, so the canonical line is the first match
    lines = L.v3_lines([1.0] * 10, 12.5) + list(L.extras(shape=(13, 13, 256), fnv=1, wall_ms=12.5,
                                                          device_forward_ms_warm=0.5))
    text = "\n".join(lines)
    t, shape, vals = suite.parse_run_log(text)
    assert t == "12.5" and shape == "13x13x256" and vals == "1 1 1 1 1"
    assert text.count(" ms") == 1


def test_parse_reference_style_logs():
    v1 = "  [Input] Dimensions: H=227, W=227, C=3\nAlexNet Serial Forward Pass completed in 742 ms\n" \
         "Final Output (first 10 values): 0.1 0.2 0.3 0.4 0.5 0.6 0.7 0.8 0.9 1...\n"
    assert suite.parse_run_log(v1) == ("742", "227x227x3", "0.1 0.2 0.3 0.4 0.5")   # the reference's mis-parse
    v2 = "shape: 13x13x256\nSample values: 44.4152 42.4612 40.6967 40.6967 40.6967\nExecution Time: 286.9 ms\n"
    assert suite.parse_run_log(v2) == ("286.9", "13x13x256", "44.4152 42.4612 40.6967 40.6967 40.6967")
    v4 = "Final Output Shape: 13x13x256\nFinal Output (first 10 values): 29.2932 25.9153 23.3255 23.3255 " \
         "23.3255 23.3255 23.3255 23.3255 23.3255 23.3255\nAlexNet MPI+CUDA Forward Pass completed in 183.2 ms\n"
    assert suite.parse_run_log(v4) == ("183.2", "13x13x256", "29.2932 25.9153 23.3255 23.3255 23.3255")


def test_failure_classification():
    assert suite.classify_failure(1, "CUDA error: no CUDA-capable device is detected")[0] == 2
    assert suite.classify_failure(1, "There are not enough slots available in the system")[0] == 3
    assert suite.classify_failure(139, "")[0] == 4
    assert suite.classify_failure(124, "")[0] == 124
    assert suite.classify_failure(7, "boom")[0] == 1


def test_csv_header_is_the_references():
    assert len(suite.CSV_HEADER) == 20
    assert suite.CSV_HEADER[4:6] == ["ProjectVariant", "NumProcesses"]
    assert suite.CSV_HEADER[-3:] == ["ExecutionTime_ms", "OutputShape", "OutputFirst5Values"]


def test_suite_end_to_end_cpu_variants(tmp_path):
    """make + run + parse for V1 and V2.2 np2 through the launchers and the mpirun shim."""
    cases = [c for c in suite.default_cases((2,)) if c.name in ("V1 Serial", "V2 2.2-scatter-halo")]
    csv_path, rows = suite.run_suite("pytest", cases, timeout=300, logs_root=str(tmp_path))
    assert [r["OverallStatusSymbol"] for r in rows] == ["✔", "✔"], rows
    assert rows[0]["OutputShape"] == "13x13x256"          # extra Final Output Shape line fixes V1
    assert rows[1]["OutputFirst5Values"].startswith("44.4152 42.4612 40.6967")
    assert os.path.exists(csv_path)


def test_mpirun_fail_stop(tmp_path):
    """One rank exiting non-zero takes the job down with its exit code (MPI_Abort semantics)."""
    prog = tmp_path / "p.py"
    prog.write_text("import os,sys,time\nr=int(os.environ['OMPI_COMM_WORLD_RANK'])\n"
                    "sys.exit(3) if r==1 else time.sleep(30)\n")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bin", "mpirun"), "--oversubscribe", "-np", "3",
                        sys.executable, str(prog)], capture_output=True, text=True, timeout=25)
    assert p.returncode == 3


def test_multi_machine_dry_run_plan():
    from b200cnn.harness import multi_machine as mm
    hosts = mm.parse_hosts("alice@nodeA:100a:1,bob@nodeB:86:1")
    assert [h.alias for h in hosts] == ["alice@nodeA", "bob@nodeB"] and hosts[1].arch == "86"
    assert mm.combined_arch_flags(hosts) == "-gencode arch=compute_100a,code=sm_100a -gencode arch=compute_86,code=sm_86"
    o = mm.Orchestrator(hosts, dry=True, shared_fs=False, verbose_rsync=False, iface="eth0")
    o.setup()
    o.run(max_np=4)
    joined = "\n".join(o.cmds)
    assert "rsync -az --delete --checksum --exclude .git/" in joined and "bob@nodeB" in joined
    # np=2 over two 1-slot hosts: one torchrun agent per host, worker started over ssh
    assert "--nnodes 2 --node-rank 1" in joined and "ssh bob@nodeB" in joined
    assert "GLOO_SOCKET_IFNAME=eth0" in joined


def test_multi_machine_local_transport_runs_two_nodes(tmp_path):
    """The multi-host path for real: two host entries = two torchrun agents on this box (nnodes 2, one rank each),
    V2.1 and V2.2 over gloo across them, workers started before the master, per-node logs, the 20-column CSV."""
    from b200cnn.harness import multi_machine as mm
    rc = mm.main(["--hosts", "nodeA:100a:1,nodeB:100a:1", "--transport", "local", "--multi-host-only", "--np", "1,2",
                  "--variants", "2.2", "--runner-args", "--height 63 --width 63", "--logs-root", str(tmp_path),
                  "--timeout", "300"])
    assert rc == 0
    sess = [d for d in os.listdir(tmp_path) if d.startswith("fullsuite_session_")]
    assert len(sess) == 1
    d = tmp_path / sess[0]
    names = os.listdir(d)
    assert "mpi_cluster_hosts.txt" in names and "main_orchestration.log" in names
    assert "run_v2_2_2_scatter_halo_np2.log" in names and "run_v2_2_2_scatter_halo_np2_node1.log" in names
    import csv
    rows = list(csv.DictReader(open(d / [n for n in names if n.startswith("summary_report_")][0])))
    assert [(r["ProjectVariant"], r["NumProcesses"]) for r in rows] == [("V2 2.2-scatter-halo", "1"), ("V2 2.2-scatter-halo", "2")]
    assert all(r["OverallStatusSymbol"] == "✔" and r["OutputShape"] == "2x2x256" for r in rows)
    assert len({r["OutputFirst5Values"] for r in rows}) == 1          # np 2 over two agents == np 1
    assert "--nnodes 2 --node-rank 1" in open(d / "run_v2_2_2_scatter_halo_np2_node1.log").read()


def test_hw_toolchain_test_and_package(tmp_path):
    from b200cnn.harness import hw
    assert hw.test(1, sizes=[128], counts=[1, 2, 3], timeout=120) == 0      # np=3 skipped (128 % 3)


def test_bash_library_functions(tmp_path):
    """scripts/common_test_utils.sh: the reference's sourceable function names (setup_test_case,
    execute_single_test_case, _log_to_csv, _print_summary_table) on top of the Python harness."""
    script = f'''
set -e
source "{ROOT}/scripts/common_test_utils.sh"
LOGS_BASE_DIR="{tmp_path}"
init_test_session unit
_detect_and_set_cuda_arch_flags
setup_test_case "V1 Serial" 1 v1_serial v1_serial && execute_single_test_case
setup_test_case "V9 Missing" 2 v9_nope v9 || true
_log_to_csv "Manual" 1 m.log true ok r.log true false ok true ok OK fine 1.5 1x1x1 "1 2 3"
_print_summary_table
echo "CSV=$CSV_SUMMARY_FILE FLAGS=$DETECTED_GPU_ARCH_FLAGS"
'''
    p = subprocess.run(["bash", "-c", script], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout + p.stderr
    csv_path = re.search(r"CSV=(\S+)", p.stdout).group(1)
    assert "compute_100a" in p.stdout or "sm_" in p.stdout
    import csv as _csv
    rows = list(_csv.DictReader(open(csv_path)))
    assert [r["ProjectVariant"] for r in rows] == ["V1 Serial", "V9 Missing", "Manual"]
    assert rows[0]["OverallStatusSymbol"] == "✔" and rows[0]["OutputShape"] == "13x13x256"
    assert rows[1]["BuildMessage"] == "Dir not found" and rows[1]["OverallStatusSymbol"] == "✘"
    assert rows[2]["ExecutionTime_ms"] == "1.5" and rows[2]["OutputFirst5Values"] == "1 2 3"
    assert os.path.exists(os.path.join(os.path.dirname(csv_path), "run_v1_serial_np1.log"))


def test_hw_native_skeleton_builds_and_passes_under_mpirun(monkeypatch, tmp_path):
    """`scaffold --native` lays down a C skeleton that compiles to ./template and passes the HW test matrix
    (counterpart of ref templates/template.c.template, which needs an MPI the image does not have)."""
    import shutil
    from b200cnn.harness import hw
    d = hw.hw_dir(97)
    shutil.rmtree(d, ignore_errors=True)
    try:
        assert hw.scaffold(97, native=True) == 0
        assert os.path.exists(os.path.join(d, "src", "template.c"))
        assert hw.test(97, sizes=[128, 256], counts=[1, 2, 4], timeout=60) == 0
    finally:
        shutil.rmtree(d, ignore_errors=True)
