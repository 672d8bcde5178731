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
executing it (used by the tests; this image has no second host).
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
    def __init__(self, hosts: List[Host], dry: bool, shared_fs: bool, verbose_rsync: bool, iface: str):
        self.hosts, self.dry, self.shared_fs, self.verbose_rsync, self.iface = hosts, dry, shared_fs, verbose_rsync, iface
        ts = datetime.datetime.now().strftime("%Y%m%d_%H%M%S")
        self.session = f"fullsuite_{ts}_{socket.gethostname().split('.')[0]}"
        self.dir = os.path.join(REPO_ROOT, "final_project", "logs", self.session)
        self.cmds: List[str] = []
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

    def setup(self):
        self.log("--- Phase 1: passwordless SSH ---")
        for h in self.hosts[1:]:
            test = f"ssh -o PasswordAuthentication=no -o BatchMode=yes -o ConnectTimeout=5 {h.alias} exit"
            if self.sh(test, check=False) != 0:
                self.sh(f"ssh-copy-id {h.alias}")
                self.sh(test)
        self.log(f"--- Phase 2: combined arch flags: {combined_arch_flags(self.hosts)} ---")
        self.log("--- Phase 3: file sync ---")
        if self.shared_fs:
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

    def launch_multi_host(self, module: str, np_: int, args: str = "") -> None:
        """np_ ranks spread over the hosts in hostfile order, one torchrun agent per host."""
        master = self.hosts[0].host
        remaining, node = np_, 0
        placed = []
        for h in self.hosts:
            k = min(h.slots, remaining)
            if k > 0:
                placed.append((h, k))
                remaining -= k
        env = f"GLOO_SOCKET_IFNAME={self.iface} " if self.iface else ""
        for node, (h, k) in enumerate(placed):
            # torchrun pins every rank to ONE OpenMP thread unless told otherwise: give each its share of the host
            omp = f"OMP_NUM_THREADS=$(( ($(nproc) + {k} - 1) / {k} )) "
            cmd = (f"{env}{omp}python -m torch.distributed.run --nnodes {len(placed)} --node-rank {node} "
                   f"--nproc-per-node {k} --master-addr {master} --master-port 29555 -m {module} {args}")
            full = cmd if node == 0 else f"ssh {h.alias} 'cd {shlex.quote(REPO_ROOT)} && {cmd}'"
            self.sh(full + (" &" if node else ""), check=False)

    def run(self, max_np: int):
        total = sum(h.slots for h in self.hosts)
        nps = [n for n in (1, 2, 4, 8) if n <= min(total, max_np)]
        self.log(f"--- suite: np in {nps} (total slots {total}) ---")
        self.sh(f"cd {shlex.quote(REPO_ROOT)} && python -m b200cnn.harness.suite --session-prefix fullsuite "
                f"--only 'V1 Serial,V3' --np 1")
        for n in nps:
            for mod in ("b200cnn.runners.v2_1", "b200cnn.runners.v2_2"):
                self.launch_multi_host(mod, n)
        self.sh(f"cd {shlex.quote(REPO_ROOT)} && python -m b200cnn.harness.suite --session-prefix fullsuite "
                f"--only 'V4,V5' --np {','.join(str(n) for n in nps if n <= self.hosts[0].slots) or '1'}")


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--hosts", default=os.environ.get("B200_HOSTS", ""))
    ap.add_argument("--config", default=None, help="JSON: {hosts:[{user,host,arch,slots}], shared_fs, iface}")
    ap.add_argument("--dry-run", action="store_true")
    ap.add_argument("--max-np", type=int, default=8)
    ns = ap.parse_args(argv)
    shared = os.environ.get("USE_SHARED_FILESYSTEM", "false") == "true"
    iface = os.environ.get("MPI_NETWORK_INTERFACE", "")
    if ns.config:
        cfg = json.load(open(ns.config))
        hosts = [Host(**h) for h in cfg["hosts"]]
        shared, iface = cfg.get("shared_fs", shared), cfg.get("iface", iface)
    else:
        hosts = parse_hosts(ns.hosts) if ns.hosts else [Host(os.environ.get("USER", ""), socket.gethostname(), "100a", 1)]
    o = Orchestrator(hosts, ns.dry_run, shared, os.environ.get("VERBOSE_RSYNC", "false") == "true", iface)
    o.setup()
    o.run(ns.max_np)
    return 0


if __name__ == "__main__":
    sys.exit(main())
