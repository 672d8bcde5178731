"""Writes analysis.ipynb (nbformat 4 JSON; no Jupyter needed to generate it): the reference notebook's flow
(ref analysis.ipynb cells 1-18: ingest -> inspect -> run statistics -> speed-up / efficiency -> exports -> plots ->
focused comparisons -> all figures) on top of b200cnn.harness.log_analysis, so each cell is a few lines.
    python scripts/make_notebook.py [analysis.ipynb]
tests/test_log_analysis.py executes every code cell in order with plain exec()."""
import json
import sys

CELLS = [
    ("md", "# Performance analysis of the five versions\n"
           "Counterpart of the reference's `analysis.ipynb`: the same warehouse tables and views "
           "(`file_index`, `summary_runs`, `run_logs`, `perf_runs`, `best_runs`, `run_stats`, `speedup`, `efficiency`), "
           "driven through `b200cnn.harness.log_analysis`. Figures are SVG (no matplotlib on the image)."),
    ("code", "# Cell 1: setup\n"
             "import os, sys\nfrom pathlib import Path\nimport pandas as pd\n"
             "ROOT = Path(os.environ.get('B200_ANALYSIS_ROOT', '.')).resolve()\n"
             "sys.path.insert(0, str(ROOT))\n"
             "from b200cnn.harness import log_analysis as la\n"
             "OUT = Path(os.environ.get('B200_ANALYSIS_OUT', ROOT / 'analysis_exports' / 'notebook'))\n"
             "OUT.mkdir(parents=True, exist_ok=True)\n"
             "DB = OUT / 'warehouse.sqlite'\n"
             "pd.set_option('display.width', 200)"),
    ("code", "# Cells 3-5: schema + ingestion (summary CSVs in all three schemas, run logs, sources)\n"
             "if DB.exists():\n    DB.unlink()\n"
             "wh = la.Warehouse(DB)\n"
             "counts = wh.ingest(Path(os.environ.get('B200_ANALYSIS_LOGS', ROOT)))\n"
             "counts"),
    ("code", "# Cell 6: database inspection\n"
             "tables = wh.df(\"SELECT name, type FROM sqlite_master WHERE type IN ('table', 'view') ORDER BY type, name\")\n"
             "per_version = wh.df('SELECT version, np, COUNT(*) AS runs, MIN(total_time_s) AS best_s FROM perf_runs GROUP BY version, np ORDER BY version, np')\n"
             "tables, per_version"),
    ("code", "# Cell 7: run statistics (n, mean, median, sd, CV per version and np; canonical version names)\n"
             "stats = wh.run_stats(canonical=True).sort_values(['version', 'np'])\n"
             "stats"),
    ("code", "# Cell 8: speed-up and efficiency, from the best and from the median time; the notebook's SQL views\n"
             "sp_best = wh.speedup('best', canonical=True)\n"
             "sp_median = wh.speedup('median', canonical=True)\n"
             "views = {v: wh.df(f'SELECT * FROM {v}') for v in ('speedup', 'efficiency')}\n"
             "sp_best, sp_median"),
    ("code", "# Cells 9-10: exports (CSV, Markdown, Parquet)\n"
             "for view, name in (('best_runs', 'best_runs.csv'), ('best_runs', 'best_runs.md'), ('run_stats', 'stats.csv'),\n"
             "                   ('perf_runs', 'all_runs.csv'), ('run_stats', 'stats.parquet'), ('radar', 'radar_table.csv')):\n"
             "    la.main(['--db', str(DB), 'export', view, str(OUT / name)])"),
    ("code", "# Cells 11-16 and 18: every figure (speed-up per version, efficiency, runtime CV, radar, LOC vs performance)\n"
             "la.main(['--db', str(DB), 'plot', 'speedup', 'efficiency', 'cv', 'radar', 'locperf', '--outdir', str(OUT / 'plots')])\n"
             "sorted(p.name for p in (OUT / 'plots').glob('*.svg'))"),
    ("code", "# Cell 17: focused comparisons - scorecard at the largest np, LOC vs best time, the radar table\n"
             "score = stats.merge(sp_median[['version', 'np', 'S', 'E']], on=['version', 'np'], how='left')\n"
             "top = score[score.np == score.np.max()]\n"
             "loc = la.loc_vs_perf(wh)\n"
             "radar = la.radar_table(wh)\n"
             "top, loc, radar"),
]


def notebook():
    cells = []
    for kind, src in CELLS:
        lines = src.split("\n")
        source = [l + "\n" for l in lines[:-1]] + [lines[-1]]
        if kind == "md":
            cells.append({"cell_type": "markdown", "metadata": {}, "source": source})
        else:
            cells.append({"cell_type": "code", "execution_count": None, "metadata": {}, "outputs": [], "source": source})
    return {"cells": cells, "metadata": {"kernelspec": {"display_name": "Python 3", "language": "python", "name": "python3"},
                                        "language_info": {"name": "python"}}, "nbformat": 4, "nbformat_minor": 5}


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 else "analysis.ipynb"
    with open(out, "w") as f:
        json.dump(notebook(), f, indent=1)
        f.write("\n")
    print(f"wrote {out} ({len(CELLS)} cells)")
