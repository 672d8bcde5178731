"""Log warehouse + analysis CLI (the L6 layer).

Same commands as the reference's ``log_analysis.py`` (ref log_analysis.py:75-292) — ``ingest``,
``query``, ``stats``, ``speedup``, ``plot``, ``export`` — plus ``scorecard`` (the notebook's median /
CV table, ref analysis.md:2137-2630).  DuckDB and matplotlib are not on this image, so the
warehouse is SQLite (stdlib) + pandas, and plots are written as dependency-free SVG.

Accepted inputs (ref log_analysis.py:54-66, analysis.md:134-142):
  * session CSVs with ``EntryTimestamp,ProjectVariant,NumProcesses`` + ``ExecutionTime_ms|Time_ms``;
  * legacy ``Timestamp,Version,NP,Time_ms`` and lower-case ``timestamp,version,np,time_ms``;
  * ``run_*.log`` files: first match of the time patterns below (our runners also print
    ``ExecutionTime_ms: <num>`` so this table is populated, unlike the reference's);
  * ``*.cpp|*.cu|*.hpp|*.h|*.cuh|*.py`` for source statistics.
Speed-up baseline is the version literally named ``V1 Serial`` at np=1 (ref :218,235).
"""
from __future__ import annotations

import argparse
import hashlib
import os
import re
import sqlite3
import sys
from datetime import datetime, timezone
from pathlib import Path
from typing import Optional

import pandas as pd

DEFAULT_DB = Path(".warehouse/cluster_logs.sqlite")
PLOTS_DIR = Path("plots")

# run-log time patterns, most specific first (ref analysis.md:311-322 lists ten variants)
TIME_PATTERNS = [
    (r"(?:Time|ExecutionTime)_ms[=:]\s*([\d.]+)", 1e-3),
    (r"AlexNet MPI\+CUDA(?:-Aware)? Forward Pass completed in ([\d.]+) ms", 1e-3),
    (r"AlexNet CUDA Forward Pass completed in ([\d.]+) ms", 1e-3),
    (r"AlexNet Serial Forward Pass completed in ([\d.]+) ms", 1e-3),
    (r"Execution Time:\s*([\d.]+) ms", 1e-3),
    (r"completed in ([\d.]+) ms", 1e-3),
    (r"Total PARALLEL time:\s*([\d.]+) seconds", 1.0),
    (r"Parallel computation finished in ([\d.]+) seconds", 1.0),
    (r"elapsed[=:]\s*([\d.]+)\s*s\b", 1.0),
    (r"([\d.]+) ms", 1e-3),
]

CANONICAL = [  # (regex on lower-cased name, canonical) — ref analysis.md:70-99
    (r"v1|serial", "V1 Serial"),
    (r"2[._]1|broadcast", "V2.1 BroadcastAll"),
    (r"2[._]2|scatter", "V2.2 ScatterHalo"),
    (r"v5|aware", "V5 MPI+CUDA-Aware"),
    (r"v4|mpi\+cuda|mpi_cuda", "V4 MPI+CUDA"),
    (r"v3|cuda", "V3 CUDA"),
]


def canonical_version_name(v: Optional[str]) -> Optional[str]:
    if v is None:
        return None
    low = str(v).lower()
    for pat, name in CANONICAL:
        if re.search(pat, low):
            return name
    return str(v)


def _sha1(p: Path) -> str:
    h = hashlib.sha1()
    with p.open("rb") as fh:
        for blk in iter(lambda: fh.read(1 << 16), b""):
            h.update(blk)
    return h.hexdigest()


def normalise_summary(df: pd.DataFrame) -> pd.DataFrame:
    cols = set(df.columns)
    out = pd.DataFrame()
    if {"EntryTimestamp", "ProjectVariant", "NumProcesses"} <= cols:
        tcol = "ExecutionTime_ms" if "ExecutionTime_ms" in cols else "Time_ms"
        if tcol not in cols:
            return out
        out = pd.DataFrame({"ts": df["EntryTimestamp"], "version": df["ProjectVariant"],
                            "np": df["NumProcesses"], "ms": df[tcol]})
    elif {"Timestamp", "Version", "NP", "Time_ms"} <= cols:
        out = pd.DataFrame({"ts": df["Timestamp"], "version": df["Version"], "np": df["NP"],
                            "ms": df["Time_ms"]})
    elif {"timestamp", "version", "np", "time_ms"} <= cols:
        out = pd.DataFrame({"ts": df["timestamp"], "version": df["version"], "np": df["np"],
                            "ms": df["time_ms"]})
    else:
        return out
    out["ts"] = pd.to_datetime(out["ts"], utc=True, errors="coerce").astype(str)
    out["np"] = pd.to_numeric(out["np"], errors="coerce")
    out["total_time_s"] = pd.to_numeric(out["ms"], errors="coerce") / 1000.0
    return out.dropna(subset=["np"])[["ts", "version", "np", "total_time_s"]]


class Warehouse:
    def __init__(self, path: Path = DEFAULT_DB):
        self.path = Path(path)
        self.path.parent.mkdir(parents=True, exist_ok=True)
        self.con = sqlite3.connect(str(self.path))
        c = self.con
        c.execute("CREATE TABLE IF NOT EXISTS file_index(relpath TEXT PRIMARY KEY, sha1 TEXT, size INT, mtime TEXT)")
        c.execute("CREATE TABLE IF NOT EXISTS summary_runs(src TEXT, ts TEXT, version TEXT, np INT, total_time_s REAL)")
        c.execute("CREATE TABLE IF NOT EXISTS run_logs(relpath TEXT, ts TEXT, version TEXT, np INT, total_time_s REAL)")
        c.execute("CREATE TABLE IF NOT EXISTS source_stats(relpath TEXT PRIMARY KEY, loc INT, func_cnt INT, include_cnt INT, cuda_kernel_cnt INT)")
        c.executescript("""
            DROP VIEW IF EXISTS efficiency; DROP VIEW IF EXISTS speedup;
            DROP VIEW IF EXISTS perf_runs; DROP VIEW IF EXISTS best_runs;
            CREATE VIEW perf_runs AS
              SELECT ts, version, np, total_time_s FROM summary_runs WHERE total_time_s IS NOT NULL
              UNION ALL
              SELECT ts, version, np, total_time_s FROM run_logs WHERE total_time_s IS NOT NULL;
            CREATE VIEW best_runs AS
              SELECT version, np, MIN(total_time_s) AS total_time_s FROM perf_runs GROUP BY version, np;
            -- per-version strong scaling, as the reference's notebook defines it (ref analysis.md:222-240):
            -- S = best time of the SAME version at np = 1 / best time at np;  E = S / np
            CREATE VIEW speedup AS
              SELECT b.version AS version, b.np AS np, base.t1 / b.total_time_s AS S
              FROM best_runs b JOIN (SELECT version, MIN(total_time_s) AS t1 FROM best_runs WHERE np = 1 GROUP BY version) base
                   ON base.version = b.version
              WHERE b.total_time_s > 1e-9;
            CREATE VIEW efficiency AS SELECT version, np, S / np AS E FROM speedup WHERE np > 0;
        """)
        c.commit()

    def df(self, sql: str) -> pd.DataFrame:
        return pd.read_sql_query(sql, self.con)

    def run_stats(self, canonical: bool = False) -> pd.DataFrame:
        """mean, sample sd, 95 % CI, median, CV per (version, np) (ref run_stats view + scorecard).
        ``canonical``: map version labels through :func:`canonical_version_name` first, so that harness CSV
        rows ("V2 2.2-scatter-halo") and run-log rows ("v2_2_2_scatter_halo_np4") pool into one group, as the
        reference's notebook does (ref analysis.md:70-99)."""
        d = self.df("SELECT version, np, total_time_s FROM perf_runs")
        if canonical and not d.empty:
            d["version"] = d.version.map(canonical_version_name)
        if d.empty:
            return pd.DataFrame(columns=["version", "np", "n", "mean_s", "sd_s", "ci95_s", "median_s", "cv"])
        g = d.groupby(["version", "np"])["total_time_s"]
        out = g.agg(n="count", mean_s="mean", sd_s="std", median_s="median").reset_index()
        out["ci95_s"] = 1.96 * out["sd_s"] / out["n"].pow(0.5)
        out["cv"] = out["sd_s"] / out["mean_s"]
        return out[["version", "np", "n", "mean_s", "sd_s", "ci95_s", "median_s", "cv"]]

    def speedup(self, stat: str = "best", canonical: bool = False) -> pd.DataFrame:
        """S = T1(V1 Serial, np=1) / T(version, np); E = S / np (ref :213-222)."""
        if stat == "best":
            b = self.df("SELECT version, np, total_time_s AS t FROM best_runs")
            if canonical and not b.empty:
                b["version"] = b.version.map(canonical_version_name)
                b = b.groupby(["version", "np"], as_index=False)["t"].min()
        else:
            b = self.run_stats(canonical).rename(columns={"median_s": "t"})[["version", "np", "t"]]
        base = b[(b.version == "V1 Serial") & (b.np == 1)]
        if base.empty:
            raise SystemExit("no 'V1 Serial' np=1 run in the warehouse: speed-up has no baseline")
        t1 = float(base.t.min())
        b["S"] = t1 / b.t
        b["E"] = b.S / b.np
        return b.sort_values("S", ascending=False).reset_index(drop=True)

    # ---- ETL -----------------------------------------------------------------------------------
    def ingest(self, root: Path, verbose: bool = False) -> dict:
        seen = dict(self.con.execute("SELECT relpath, sha1 FROM file_index").fetchall())
        n = dict(files=0, summary_rows=0, run_logs=0, sources=0)
        for p in sorted(Path(root).rglob("*")):
            if p.is_dir() or ".git" in p.parts or ".warehouse" in p.parts or \
                    p.suffix.lower() in (".sqlite", ".sqlite-journal", ".duckdb"):
                continue
            rel = str(p.relative_to(root))
            try:
                sha = _sha1(p)
            except OSError:
                continue
            if seen.get(rel) == sha:
                continue                      # SHA-1 dedup: unchanged files are not re-ingested
            st = p.stat()
            mtime = datetime.fromtimestamp(st.st_mtime, timezone.utc).isoformat()
            suf = p.suffix.lower()
            if suf == ".csv":
                try:
                    nd = normalise_summary(pd.read_csv(p))
                except Exception:
                    nd = pd.DataFrame()
                if not nd.empty:
                    self.con.execute("DELETE FROM summary_runs WHERE src=?", (rel,))
                    self.con.executemany("INSERT INTO summary_runs VALUES (?,?,?,?,?)",
                                         [(rel, r.ts, r.version, int(r.np),
                                           None if pd.isna(r.total_time_s) else float(r.total_time_s))
                                          for r in nd.itertuples()])
                    n["summary_rows"] += len(nd)
            elif suf == ".log" and "run_" in p.name:
                txt = p.read_text(errors="ignore")
                sec = None
                for pat, scale in TIME_PATTERNS[:1]:       # the reference only honours the first form
                    m = re.search(pat, txt)
                    if m:
                        sec = float(m.group(1)) * scale
                if sec is not None:
                    vm = re.search(r"v\d(?:_\d[._]\d[_\w]*)?", rel)
                    nm = re.search(r"np(\d+)", rel)
                    self.con.execute("DELETE FROM run_logs WHERE relpath=?", (rel,))
                    self.con.execute("INSERT INTO run_logs VALUES (?,?,?,?,?)",
                                     (rel, mtime, vm.group(0) if vm else None,
                                      int(nm.group(1)) if nm else None, sec))
                    n["run_logs"] += 1
            elif suf in (".cpp", ".cu", ".hpp", ".h", ".cuh", ".py", ".c"):
                code = p.read_text(errors="ignore")
                self.con.execute("INSERT OR REPLACE INTO source_stats VALUES (?,?,?,?,?)",
                                 (rel, code.count("\n"),
                                  len(re.findall(r"\b[A-Za-z_]\w*\s*\([^)]*\)\s*\{", code)) +
                                  len(re.findall(r"^\s*def \w+", code, re.M)),
                                  code.count("#include") + len(re.findall(r"^\s*(?:import|from) ", code, re.M)),
                                  code.count("__global__")))
                n["sources"] += 1
            self.con.execute("INSERT OR REPLACE INTO file_index VALUES (?,?,?,?)", (rel, sha, st.st_size, mtime))
            n["files"] += 1
            if verbose:
                print(f"indexed {rel}")
        self.con.commit()
        return n


def loc_vs_perf(wh: "Warehouse") -> pd.DataFrame:
    src = wh.df("SELECT relpath, loc, cuda_kernel_cnt FROM source_stats")
    if src.empty:
        return pd.DataFrame(columns=["version", "loc", "kernels", "best_s"])
    src["version"] = src.relpath.map(lambda r: canonical_version_name(r) if re.search(r"v\d|serial|cuda|mpi", r.lower()) else None)
    loc = src.dropna(subset=["version"]).groupby("version").agg(loc=("loc", "sum"), kernels=("cuda_kernel_cnt", "sum")).reset_index()
    best = wh.df("SELECT version, MIN(total_time_s) AS best_s FROM perf_runs GROUP BY version")
    best["version"] = best.version.map(canonical_version_name)
    best = best.groupby("version").best_s.min().reset_index()
    return loc.merge(best, on="version", how="outer").sort_values("version")


def time_from_log(text: str) -> Optional[float]:
    """Seconds from a run log using the full pattern ladder (ref analysis.md:311-322)."""
    for pat, scale in TIME_PATTERNS:
        m = re.search(pat, text)
        if m:
            return float(m.group(1)) * scale
    return None


# ---- dependency-free SVG charts ------------------------------------------------------------------
def _svg_axes(w, h, title, xlabel, ylabel):
    return [f'<svg xmlns="http://www.w3.org/2000/svg" width="{w}" height="{h}" font-family="sans-serif" font-size="12">',
            f'<rect width="{w}" height="{h}" fill="white"/>',
            f'<text x="{w/2}" y="18" text-anchor="middle" font-size="14">{title}</text>',
            f'<text x="{w/2}" y="{h-6}" text-anchor="middle">{xlabel}</text>',
            f'<text x="14" y="{h/2}" text-anchor="middle" transform="rotate(-90 14 {h/2})">{ylabel}</text>',
            f'<line x1="50" y1="{h-40}" x2="{w-15}" y2="{h-40}" stroke="black"/>',
            f'<line x1="50" y1="30" x2="50" y2="{h-40}" stroke="black"/>']


def svg_line(path: Path, xs, ys, title, xlabel, ylabel):
    w, h = 480, 320
    xmax, ymax = max(max(xs), 1), max(max(ys) * 1.1, 1e-9)
    px = lambda x: 50 + (w - 65) * x / xmax
    py = lambda y: (h - 40) - (h - 70) * y / ymax
    el = _svg_axes(w, h, title, xlabel, ylabel)
    el.append('<polyline fill="none" stroke="#1f77b4" stroke-width="2" points="' +
              " ".join(f"{px(x):.1f},{py(y):.1f}" for x, y in zip(xs, ys)) + '"/>')
    for x, y in zip(xs, ys):
        el.append(f'<circle cx="{px(x):.1f}" cy="{py(y):.1f}" r="3.5" fill="#1f77b4"/>')
        el.append(f'<text x="{px(x):.1f}" y="{h-26}" text-anchor="middle">{x:g}</text>')
        el.append(f'<text x="{px(x)+6:.1f}" y="{py(y)-6:.1f}">{y:.2f}</text>')
    el.append("</svg>")
    path.write_text("\n".join(el))


def svg_bars(path: Path, labels, vals, title, ylabel, ymax=1.1):
    w, h = max(480, 60 + 46 * len(labels)), 340
    el = _svg_axes(w, h, title, "", ylabel)
    bw = (w - 80) / max(len(labels), 1)
    for i, (lab, v) in enumerate(zip(labels, vals)):
        x = 55 + i * bw
        bh = (h - 90) * min(v, ymax) / ymax
        el.append(f'<rect x="{x:.1f}" y="{h-60-bh:.1f}" width="{bw*0.8:.1f}" height="{bh:.1f}" fill="#2ca02c"/>')
        el.append(f'<text x="{x+bw*0.4:.1f}" y="{h-64-bh:.1f}" text-anchor="middle" font-size="10">{v:.2f}</text>')
        el.append(f'<text x="{x+bw*0.4:.1f}" y="{h-46}" text-anchor="end" font-size="9" '
                  f'transform="rotate(-35 {x+bw*0.4:.1f} {h-46})">{lab}</text>')
    el.append("</svg>")
    path.write_text("\n".join(el))


def svg_grouped_bars(path: Path, groups, series, values, title, ylabel, notes=None):
    """Bars of ``values[g][s]`` grouped by ``groups`` with one colour per ``series`` (the notebook's CV chart:
    version on the x axis, np as hue, n as annotation; ref analysis.md:2305-2355)."""
    colours = ["#1f77b4", "#ff7f0e", "#2ca02c", "#d62728", "#9467bd", "#8c564b"]
    w, h = max(520, 90 + 34 * len(groups) * max(len(series), 1)), 360
    vmax = max([v for g in groups for v in values[g].values() if v == v] + [1e-9]) * 1.15
    el = _svg_axes(w, h, title, "", ylabel)
    gw = (w - 80) / max(len(groups), 1)
    bw = gw * 0.8 / max(len(series), 1)
    for gi, g in enumerate(groups):
        for si, sname in enumerate(series):
            v = values[g].get(sname)
            if v is None or v != v:
                continue
            x = 55 + gi * gw + si * bw
            bh = (h - 100) * v / vmax
            el.append(f'<rect x="{x:.1f}" y="{h-70-bh:.1f}" width="{bw*0.92:.1f}" height="{bh:.1f}" fill="{colours[si % 6]}"/>')
            lab = f"{v:.2f}" + (f" (n={notes[g][sname]})" if notes and notes.get(g, {}).get(sname) else "")
            el.append(f'<text x="{x+bw*0.46:.1f}" y="{h-74-bh:.1f}" text-anchor="middle" font-size="9">{lab}</text>')
        el.append(f'<text x="{55+gi*gw+gw*0.4:.1f}" y="{h-52}" text-anchor="middle" font-size="10">{_label(str(g))}</text>')
    for si, sname in enumerate(series):
        el.append(f'<rect x="{60+si*70}" y="{h-30}" width="10" height="10" fill="{colours[si % 6]}"/>')
        el.append(f'<text x="{74+si*70}" y="{h-21}" font-size="10">np={sname}</text>')
    el.append("</svg>")
    path.write_text("\n".join(el))


def svg_radar(path: Path, axes, rows, title):
    """Radar chart: ``rows`` = {name: [value per axis in 0..1]} (ref analysis.md:2357-2460: NP1 performance, max
    speed-up, max efficiency and code volume per version, each min-max normalised over the versions)."""
    import math
    colours = ["#1f77b4", "#ff7f0e", "#2ca02c", "#d62728", "#9467bd", "#8c564b"]
    w, h, cx, cy, R = 560, 460, 250, 240, 170
    el = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{w}" height="{h}" font-family="sans-serif" font-size="11">',
          f'<rect width="{w}" height="{h}" fill="white"/>',
          f'<text x="{w/2}" y="18" text-anchor="middle" font-size="14">{title}</text>']
    k = len(axes)
    pt = lambda i, r: (cx + R * r * math.sin(2 * math.pi * i / k), cy - R * r * math.cos(2 * math.pi * i / k))
    for ring in (0.25, 0.5, 0.75, 1.0):
        el.append('<polygon fill="none" stroke="#ccc" points="' + " ".join("%.1f,%.1f" % pt(i, ring) for i in range(k)) + '"/>')
    for i, a in enumerate(axes):
        x, y = pt(i, 1.0)
        el.append(f'<line x1="{cx}" y1="{cy}" x2="{x:.1f}" y2="{y:.1f}" stroke="#bbb"/>')
        lx, ly = pt(i, 1.13)
        el.append(f'<text x="{lx:.1f}" y="{ly:.1f}" text-anchor="middle">{a}</text>')
    for ri, (name, vals) in enumerate(rows.items()):
        c = colours[ri % 6]
        el.append(f'<polygon fill="{c}" fill-opacity="0.15" stroke="{c}" stroke-width="2" points="' +
                  " ".join("%.1f,%.1f" % pt(i, max(0.0, min(1.0, v))) for i, v in enumerate(vals)) + '"/>')
        el.append(f'<rect x="{w-140}" y="{40+ri*18}" width="10" height="10" fill="{c}"/>')
        el.append(f'<text x="{w-125}" y="{49+ri*18}">{name}</text>')
    el.append("</svg>")
    path.write_text("\n".join(el))


def svg_scatter(path: Path, pts, title, xlabel, ylabel):
    """Labelled scatter (LOC vs best time per version; ref analysis.md 'LOC vs performance')."""
    w, h = 520, 340
    xmax = max([p[0] for p in pts] + [1]) * 1.1
    ymax = max([p[1] for p in pts] + [1e-9]) * 1.15
    el = _svg_axes(w, h, title, xlabel, ylabel)
    for x, y, lab in pts:
        px, py = 50 + (w - 65) * x / xmax, (h - 40) - (h - 70) * y / ymax
        el.append(f'<circle cx="{px:.1f}" cy="{py:.1f}" r="4" fill="#d62728"/>')
        el.append(f'<text x="{px+6:.1f}" y="{py-5:.1f}" font-size="10">{lab} ({x:g} LOC, {y:.3g} s)</text>')
    el.append("</svg>")
    path.write_text("\n".join(el))


def radar_table(wh: "Warehouse") -> pd.DataFrame:
    """Per canonical version: NP1 performance (1 / median time at np = 1), max speed-up and max efficiency over np
    (median based, baseline V1 Serial), log10 of the lines of code; raw values and min-max normalised columns."""
    import math
    st = wh.run_stats(canonical=True)
    if st.empty:
        return pd.DataFrame()
    try:
        sp = wh.speedup("median", canonical=True)
    except SystemExit:
        sp = pd.DataFrame(columns=["version", "np", "S", "E"])
    loc = loc_vs_perf(wh)
    rows = []
    for v in sorted(st.version.dropna().unique()):
        one = st[(st.version == v) & (st.np == 1)]
        s = sp[sp.version == v]
        l = loc[loc.version == v]["loc"] if not loc.empty else []
        rows.append(dict(version=v,
                         np1_perf=1.0 / float(one.median_s.iloc[0]) if len(one) and one.median_s.iloc[0] > 0 else float("nan"),
                         max_speedup=float(s.S.max()) if len(s) else float("nan"),
                         max_efficiency=float(s.E.max()) if len(s) else float("nan"),
                         log10_loc=math.log10(float(l.iloc[0])) if len(l) and l.iloc[0] and l.iloc[0] == l.iloc[0] else float("nan")))
    d = pd.DataFrame(rows)
    for c in ("np1_perf", "max_speedup", "max_efficiency", "log10_loc"):
        lo, hi = d[c].min(), d[c].max()
        d[c + "_norm"] = ((d[c] - lo) / (hi - lo)) if hi > lo else 0.5
    return d


def _label(v: str) -> str:
    return v.replace("V2 2.1-broadcast-all", "V2 Broadcast").replace("V2 2.2-scatter-halo", "V2 Scatter")


# ---- CLI -------------------------------------------------------------------------------------------
def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="log_analysis", description=__doc__.split("\n")[0])
    ap.add_argument("--db", default=str(DEFAULT_DB))
    sub = ap.add_subparsers(dest="cmd", required=True)
    a = sub.add_parser("ingest"); a.add_argument("--root", required=True); a.add_argument("--rebuild", action="store_true")
    a.add_argument("-v", "--verbose", action="store_true")
    a = sub.add_parser("query"); a.add_argument("sql")
    sub.add_parser("stats")
    a = sub.add_parser("speedup"); a.add_argument("--stat", choices=["best", "median"], default="best")
    a = sub.add_parser("scorecard")
    sub.add_parser("locperf")
    a = sub.add_parser("legacy-graphs"); a.add_argument("csv"); a.add_argument("--outdir", default=str(PLOTS_DIR))
    a = sub.add_parser("plot"); a.add_argument("plots", nargs="+", choices=["speedup", "efficiency", "cv", "radar", "locperf"])
    a.add_argument("--outdir", default=str(PLOTS_DIR))
    a = sub.add_parser("export"); a.add_argument("view"); a.add_argument("outfile")
    ns = ap.parse_args(argv)
    db = Path(ns.db)
    pd.set_option("display.width", 200); pd.set_option("display.max_rows", 500)

    if ns.cmd == "ingest":
        if ns.rebuild and db.exists():
            db.unlink()
            print("wiped existing warehouse")
        n = Warehouse(db).ingest(Path(ns.root), ns.verbose)
        print(f"ingest complete: {n}")
        return 0
    wh = Warehouse(db)
    if ns.cmd == "query":
        print(wh.df(ns.sql).to_string(index=False))
    elif ns.cmd == "stats":
        print(wh.run_stats().sort_values(["version", "np"]).to_string(index=False))
    elif ns.cmd == "speedup":
        print(wh.speedup(ns.stat).to_string(index=False))
    elif ns.cmd == "scorecard":
        st = wh.run_stats(canonical=True)
        sp = wh.speedup("median", canonical=True)[["version", "np", "S", "E"]]
        print(st.merge(sp, on=["version", "np"], how="left").sort_values(["version", "np"]).to_string(index=False))
    elif ns.cmd == "locperf":
        # lines of code per version directory vs its best time (ref analysis.md "LOC vs performance")
        print(loc_vs_perf(wh).to_string(index=False))
    elif ns.cmd == "legacy-graphs":
        # ref final_project/old_logs/generate_graphs.py: bar chart of time per (version, np) from a
        # run_summary_*.csv in any accepted schema
        nd = normalise_summary(pd.read_csv(ns.csv))
        if nd.empty:
            print("unrecognised CSV schema", file=sys.stderr)
            return 1
        g = nd.groupby(["version", "np"])["total_time_s"].min().reset_index()
        out = Path(ns.outdir); out.mkdir(parents=True, exist_ok=True)
        svg_bars(out / "legacy_times.svg", [f"{_label(v)} np={int(n)}" for v, n in zip(g.version, g.np)],
                 list(g.total_time_s), "Best time per version / np", "seconds", ymax=max(g.total_time_s) * 1.1)
        print(f"plot -> {out}/legacy_times.svg")
    elif ns.cmd == "plot":
        out = Path(ns.outdir); out.mkdir(parents=True, exist_ok=True)
        sp = wh.speedup("best")
        if "speedup" in ns.plots:
            for v in sorted(sp.version.unique()):
                s = sp[sp.version == v].sort_values("np")
                f = out / f"speedup_{v.replace('/', '-').replace(' ', '_')}.svg"
                svg_line(f, list(s.np), list(s.S), f"Speed-up - {_label(v)}", "np", "Speed-up S")
            print(f"plots -> {out}/speedup_*.svg")
        if "efficiency" in ns.plots:
            s = sp.sort_values(["version", "np"])
            svg_bars(out / "efficiency.svg", [f"{_label(v)} np={int(n)}" for v, n in zip(s.version, s.np)],
                     list(s.E), "Strong-scaling efficiency", "Efficiency E")
            print(f"plot -> {out}/efficiency.svg")
        if "cv" in ns.plots:             # runtime variability: CV = sd / mean per (version, np), n annotated
            st = wh.run_stats(canonical=True).dropna(subset=["cv"])
            groups = sorted(st.version.unique())
            series = sorted(int(n) for n in st.np.unique())
            vals = {g: {int(r.np): float(r.cv) for r in st[st.version == g].itertuples()} for g in groups}
            notes = {g: {int(r.np): int(r.n) for r in st[st.version == g].itertuples()} for g in groups}
            svg_grouped_bars(out / "runtime_cv.svg", groups, series, vals, "Runtime CV (lower = more stable)",
                             "CV = sd / mean", notes)
            print(f"plot -> {out}/runtime_cv.svg")
        if "radar" in ns.plots:
            rt = radar_table(wh)
            if not rt.empty:
                cols = ["np1_perf_norm", "max_speedup_norm", "max_efficiency_norm", "log10_loc_norm"]
                svg_radar(out / "radar.svg", ["NP1 perf", "max speed-up", "max efficiency", "code volume"],
                          {r.version: [0.0 if v != v else float(v) for v in (getattr(r, c) for c in cols)]
                           for r in rt.itertuples()}, "Versions compared (min-max normalised)")
                print(f"plot -> {out}/radar.svg")
        if "locperf" in ns.plots:
            lp = loc_vs_perf(wh).dropna(subset=["loc", "best_s"])
            if not lp.empty:
                svg_scatter(out / "loc_vs_perf.svg", [(float(r.loc), float(r.best_s), r.version) for r in lp.itertuples()],
                            "Code volume vs best time", "lines of code", "best time (s)")
                print(f"plot -> {out}/loc_vs_perf.svg")
    elif ns.cmd == "export":
        special = {"run_stats": wh.run_stats, "speedup": lambda: wh.speedup("best"), "radar": lambda: radar_table(wh)}
        try:
            df = special[ns.view]() if ns.view in special else wh.df(f"SELECT * FROM {ns.view}")
        except Exception as e:
            print(f"cannot select from '{ns.view}': {e}", file=sys.stderr)
            return 1
        of = Path(ns.outfile); of.parent.mkdir(parents=True, exist_ok=True)
        if of.suffix.lower() == ".csv":
            df.to_csv(of, index=False)
        elif of.suffix.lower() == ".parquet":
            df.to_parquet(of, index=False, engine="pyarrow")
        elif of.suffix.lower() == ".md":
            of.write_text(df.to_markdown(index=False) if hasattr(df, "to_markdown") else df.to_string(index=False))
        else:
            print("outfile must end with .csv, .parquet or .md", file=sys.stderr)
            return 1
        print(f"exported {len(df):,} rows -> {of}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
