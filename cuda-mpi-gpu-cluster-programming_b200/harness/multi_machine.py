"""Multi-machine suite (ref scripts/2_final_multi_machine.sh): passwordless-ssh check, union of
GPU arch flags, rsync of the project to the workers, hostfile, then the suite with
np <= min(total slots, 8).

Configuration (the reference edits bash variables in place, ref :26-43): a JSON file or env vars
    B200_HOSTS="user@host:arch[:slots],user@host2:arch[:slots]"   first host = master
    USE_SHARED_FILESYSTEM=true|false   VERBOSE_RSYNC=true|false   MPI_NETWORK_INTERFACE=eth0
Ranks are started on every host with ``torch.distributed.run --nnodes H --node-rank i`` over ssh
(there is no MPI daemon); the CPU variants (V1, V2.1, V2.2 over gloo) run across hosts.  The
peer-memory GPU variants (V4/V5) are limited to one NVLink domain, so across hosts only their
single-node configurations are run, on the master.  ``--dry-run`` prints every command instead of
executing it; ``--transport local`` treats every host entry as this machine (one torchrun agent per entry, no ssh), so
the whole multi-node path — rendezvous across agents, per-node logs, fail-stop, the session CSV — runs for real on
one box (this image has no second host and no ssh client).
"""
from __future__ import annotations

import argparse
import datetime
import json
import os
import shlex
import socket
import subprocess
import sys
from dataclasses import dataclass
from typing import List

from .. import REPO_ROOT


@dataclass
class Host:
    user: str
    host: str
    arch: str
    slots: int = 1

    @property
    def alias(self) -> str:
        return f"{self.user}@{self.host}" if self.user else self.host


def parse_hosts(spec: str) -> List[Host]:
    out = []
    for ent in [e.strip() for e in spec.replace(";", ",").split(",") if e.strip()]:
        parts = ent.split(":")
        uh = parts[0]
        user, host = uh.split("@", 1) if "@" in uh else (os.environ.get("USER", ""), uh)
        out.append(Host(user, host, parts[1] if len(parts) > 1 else "100a",
                        int(parts[2]) if len(parts) > 2 else 1))
    return out


def combined_arch_flags(hosts: List[Host]) -> str:
    """Union of -gencode flags (ref :243-256).  Only sm_100a builds exist here; others are listed so
    the log shows what the cluster contains."""
    flags = []
    for a in sorted({h.arch for h in hosts}):
        flags.append(f"-gencode arch=compute_{a},code=sm_{a}")
    return " ".join(flags)


class Orchestrator:
    def __init__(self, hosts: List[Host], dry: bool, shared_fs: bool, verbose_rsync: bool, iface: str,
                 transport: str = "ssh", timeout: int = 900, runner_args: str = "", logs_root: str = ""):
        self.hosts, self.dry, self.shared_fs, self.verbose_rsync, self.iface = hosts, dry, shared_fs, verbose_rsync, iface
        self.transport, self.timeout, self.runner_args = transport, timeout, runner_args
        self.ts = datetime.datetime.now().strftime("%Y%m%d_%H%M%S")
        self.session = f"fullsuite_session_{self.ts}_{socket.gethostname().split('.')[0]}"
        self.dir = os.path.join(logs_root or os.path.join(REPO_ROOT, "final_project", "logs"), self.session)
        self.csv_path = os.path.join(self.dir, f"summary_report_{self.ts}.csv")
        self.cmds: List[str] = []
        self.rows: List[dict] = []
        if not dry:
            os.makedirs(self.dir, exist_ok=True)

    def log(self, msg: str):
        line = f"[{datetime.datetime.now().strftime('%H:%M:%S')}] {msg}"
        print(line, flush=True)
        if not self.dry:
            with open(os.path.join(self.dir, "main_orchestration.log"), "a") as f:
                f.write(line + "\n")

    def sh(self, cmd: str, check: bool = True) -> int:
        self.cmds.append(cmd)
        if self.dry:
            print("  $ " + cmd)
            return 0
        rc = subprocess.run(cmd, shell=True).returncode
        if rc and check:
            self.log(f"ERROR: command failed rc={rc}: {cmd}")
            raise SystemExit(1)
        return rc

    def on_host(self, h: Host, cmd: str) -> str:
        """The shell line that runs ``cmd`` in the project directory of host ``h``.  ``local`` transport: every
        "host" is this machine (one agent process per host entry) — the multi-node rendezvous, per-node logs and
        the CSV are exercised for real without a second box."""
        inner = f"cd {shlex.quote(REPO_ROOT)} && {cmd}"
        if self.transport == "local":
            return f"bash -c {shlex.quote(inner)}"
        return f"ssh {h.alias} {shlex.quote(inner)}"

    def setup(self):
        self.log("--- Phase 1: passwordless SSH ---")
        if self.transport == "local":
            self.log("local transport: every host entry is this machine, no ssh")
        else:
            for h in self.hosts[1:]:
                test = f"ssh -o PasswordAuthentication=no -o BatchMode=yes -o ConnectTimeout=5 {h.alias} exit"
                if self.sh(test, check=False) != 0:
                    self.sh(f"ssh-copy-id {h.alias}")
                    self.sh(test)
        self.log(f"--- Phase 2: combined arch flags: {combined_arch_flags(self.hosts)} ---")
        self.log("--- Phase 3: file sync ---")
        if self.shared_fs or self.transport == "local":
            self.log("shared filesystem: skipping rsync")
        else:
            for h in self.hosts[1:]:
                self.sh(f"ssh {h.alias} mkdir -p {shlex.quote(REPO_ROOT)}")
                v = "-avz" if self.verbose_rsync else "-az"
                self.sh(f"rsync {v} --delete --checksum --exclude .git/ {shlex.quote(REPO_ROOT)}/ "
                        f"{h.alias}:{shlex.quote(REPO_ROOT)}/")
        self.log("--- Phase 4: hostfile ---")
        hf = "\n".join(f"{h.host} slots={h.slots}" for h in self.hosts) + "\n"
        self.hostfile = os.path.join(self.dir, "mpi_cluster_hosts.txt")
        if not self.dry:
            with open(self.hostfile, "w") as f:
                f.write(hf)
        print(hf, end="")

    def placement(self, np_: int):
        """np_ ranks over the hosts in hostfile order (fill a host's slots, then the next: Open MPI's by-slot)."""
        remaining, placed = np_, []
        for h in self.hosts:
            k = min(h.slots, remaining)
            if k > 0:
                placed.append((h, k))
                remaining -= k
        return placed

    def launch_multi_host(self, name: str, tag: str, module: str, np_: int) -> dict:
        """One case: a torchrun agent per host (workers first, in the background; the master's agent in the
        foreground), fail-stop, then the master's log parsed into a CSV row like every single-node case."""
        from . import suite
        placed = self.placement(np_)
        master = "127.0.0.1" if self.transport == "local" else self.hosts[0].host
        port = 29555 + (len(self.rows) % 200)
        if self.transport == "local" and not self.dry:       # one box: ask the kernel for a port nobody holds
            with socket.socket() as sk:
                sk.bind(("127.0.0.1", 0))
                port = sk.getsockname()[1]
        env = f"GLOO_SOCKET_IFNAME={self.iface} " if self.iface else ""
        lines = []
        for node, (h, k) in enumerate(placed):
            # torchrun pins every rank to ONE OpenMP thread unless told otherwise: give each its share of the host
            omp = f"OMP_NUM_THREADS=$(( ($(nproc) + {k} - 1) / {k} )) "
            cmd = (f"{env}{omp}PYTHONPATH={shlex.quote(REPO_ROOT)} {shlex.quote(sys.executable)} -m torch.distributed.run "
                   f"--nnodes {len(placed)} --node-rank {node} --nproc-per-node {k} --master-addr {master} "
                   f"--master-port {port} -m {module} {self.runner_args}").rstrip()
            lines.append(cmd if node == 0 else self.on_host(h, cmd))
        self.cmds.extend(lines)
        if self.dry:
            for ln in lines[1:]:
                print("  $ " + ln + " &")
            print("  $ " + lines[0])
            return {}
        rl = os.path.join(self.dir, f"run_{tag}_np{np_}.log")
        row = {h: "" for h in suite.CSV_HEADER}
        row.update(SessionID=self.session, MachineID=socket.gethostname(), GitCommit=suite.git_commit(),
                   EntryTimestamp=datetime.datetime.now().isoformat(timespec="seconds"), ProjectVariant=name,
                   NumProcesses=np_, MakeLogFile="N/A", BuildSucceeded="true", BuildMessage="no build step (Python runner)",
                   RunLogFile=os.path.basename(rl))
        workers = []
        for node, ln in enumerate(lines[1:], 1):
            wl = open(os.path.join(self.dir, f"run_{tag}_np{np_}_node{node}.log"), "w")
            wl.write(f"$ {ln}\n")
            wl.flush()
            workers.append((subprocess.Popen(ln, shell=True, stdout=wl, stderr=subprocess.STDOUT,
                                             start_new_session=True), wl))
        with open(rl, "w") as f:
            f.write(f"$ {lines[0]}\n")
            f.flush()
            try:
                rc = subprocess.run(lines[0], shell=True, cwd=REPO_ROOT, stdout=f, stderr=subprocess.STDOUT,
                                    timeout=self.timeout).returncode
            except subprocess.TimeoutExpired:
                f.write("\n[harness] timeout\n")
                rc = 124
        for pr, wl in workers:                       # fail-stop: a worker that outlives the master is killed
            try:
                wrc = pr.wait(timeout=30 if rc == 0 else 1)
            except subprocess.TimeoutExpired:
                try:
                    os.killpg(pr.pid, 15)
                except OSError:
                    pass
                wrc = 124
            wl.close()
            rc = rc or wrc
        text = open(rl).read()
        if rc == 0:
            row.update(RunCommandSucceeded="true", RunEnvironmentWarning="false", RunMessage="Run OK")
        else:
            code, msg = suite.classify_failure(rc, text)
            row.update(RunCommandSucceeded="false", RunEnvironmentWarning="true" if code in (2, 3) else "false",
                       RunMessage=msg)
        t, shape, vals = suite.parse_run_log(text.split("\n", 1)[1] if "\n" in text else text)
        parsed = rc == 0 and None not in (t, shape, vals)
        row.update(ParseSucceeded="true" if parsed else "false",
                   ParseMessage="Parsed OK" if parsed else "missing: " + ",".join(
                       n for n, v in (("time", t), ("shape", shape), ("values", vals)) if v is None),
                   ExecutionTime_ms=t or "", OutputShape=shape or "", OutputFirst5Values=vals or "")
        if parsed:
            row.update(OverallStatusSymbol="✔", OverallStatusMessage=f"OK ({len(placed)} hosts)")
        else:
            row.update(OverallStatusSymbol="✘", OverallStatusMessage=row["RunMessage"] if rc else row["ParseMessage"])
        self.rows.append(row)
        import csv
        new = not os.path.exists(self.csv_path)
        with open(self.csv_path, "a", newline="") as f:
            wtr = csv.DictWriter(f, fieldnames=suite.CSV_HEADER)
            if new:
                wtr.writeheader()
            wtr.writerow(row)
        self.log(f"{name} np={np_} on {len(placed)} host(s): {row['OverallStatusSymbol']} {row['ExecutionTime_ms']} ms "
                 f"{row['OutputShape']}")
        return row

    def run(self, max_np: int, single_node: bool = True, np_list=None, variants=("2.1", "2.2")):
        total = sum(h.slots for h in self.hosts)
        nps = [n for n in (np_list or (1, 2, 4, 8)) if n <= min(total, max_np)]
        self.log(f"--- suite: np in {nps} (total slots {total}) ---")
        if single_node:
            self.sh(f"cd {shlex.quote(REPO_ROOT)} && python -m b200cnn.harness.suite --session-prefix fullsuite "
                    f"--only 'V1 Serial,V3' --np 1", check=False)
        for n in nps:
            for var, name, tag, mod in (("2.1", "V2 2.1-broadcast-all", "v2_2_1_broadcast_all", "b200cnn.runners.v2_1"),
                                        ("2.2", "V2 2.2-scatter-halo", "v2_2_2_scatter_halo", "b200cnn.runners.v2_2")):
                if var in variants:
                    self.launch_multi_host(name, tag, mod, n)
        if single_node:
            self.sh(f"cd {shlex.quote(REPO_ROOT)} && python -m b200cnn.harness.suite --session-prefix fullsuite "
                    f"--only 'V4,V5' --np {','.join(str(n) for n in nps if n <= self.hosts[0].slots) or '1'}",
                    check=False)
        if self.rows:
            from . import suite
            print(suite.summary_table(self.rows))
            print(f"[harness] CSV: {self.csv_path}")


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--hosts", default=os.environ.get("B200_HOSTS", ""))
    ap.add_argument("--config", default=None, help="JSON: {hosts:[{user,host,arch,slots}], shared_fs, iface}")
    ap.add_argument("--dry-run", action="store_true")
    ap.add_argument("--max-np", type=int, default=8)
    ap.add_argument("--transport", choices=["ssh", "local"], default=os.environ.get("B200_TRANSPORT", "ssh"),
                    help="local: every host entry is this machine (one torchrun agent per entry, no ssh)")
    ap.add_argument("--multi-host-only", action="store_true",
                    help="skip the single-node parts of the suite (V1, V3, V4/V5 on the master)")
    ap.add_argument("--runner-args", default="", help="extra flags for the version runners, e.g. '--height 63 --width 63'")
    ap.add_argument("--np", dest="np_list", default="", help="comma list of process counts (default 1,2,4,8 up to the slots)")
    ap.add_argument("--variants", default="2.1,2.2", help="which multi-host versions to run")
    ap.add_argument("--timeout", type=int, default=900)
    ap.add_argument("--logs-root", default="")
    ns = ap.parse_args(argv)
    shared = os.environ.get("USE_SHARED_FILESYSTEM", "false") == "true"
    iface = os.environ.get("MPI_NETWORK_INTERFACE", "")
    if ns.config:
        cfg = json.load(open(ns.config))
        hosts = [Host(**h) for h in cfg["hosts"]]
        shared, iface = cfg.get("shared_fs", shared), cfg.get("iface", iface)
    else:
        hosts = parse_hosts(ns.hosts) if ns.hosts else [Host(os.environ.get("USER", ""), socket.gethostname(), "100a", 1)]
    o = Orchestrator(hosts, ns.dry_run, shared, os.environ.get("VERBOSE_RSYNC", "false") == "true", iface,
                     transport=ns.transport, timeout=ns.timeout, runner_args=ns.runner_args, logs_root=ns.logs_root)
    o.setup()
    o.run(ns.max_np, single_node=not ns.multi_host_only,
          np_list=[int(v) for v in ns.np_list.split(",") if v] or None, variants=tuple(ns.variants.split(",")))
    bad = [r for r in o.rows if r.get("OverallStatusSymbol") != "✔"]
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
