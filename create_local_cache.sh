#!/usr/bin/env bash
# Local build cache + pinned-environment record (role of the reference's create_local_cache.sh, which fills a local
# Nix binary cache so that a second machine builds nothing; this image has no Nix and no network, so what is worth
# caching is what THIS tree builds and the exact tool / package versions it was built with).
#   ./create_local_cache.sh             build everything, write env.lock, pack build_cache_local/b200cnn-<key>.tgz
#   ./create_local_cache.sh --restore   unpack the newest archive whose key matches this checkout's sources
#   ./create_local_cache.sh --env       only (re)write env.lock
# The key is a hash over the native sources and the toolchain versions: a stale cache is never restored.
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"; cd "$ROOT"
CACHE="$ROOT/build_cache_local"
key() {
  { cat csrc/cpu/*.cpp csrc/cuda/*.cu csrc/cuda/*.cuh include/*.h baseline/mpishim/mpishim.c 2>/dev/null
    nvcc --version 2>/dev/null | tail -1; g++ --version | head -1; } | sha1sum | cut -c1-12
}
write_env() {
  { echo "# pinned environment of this checkout (counterpart of the reference's shell.nix pin: nixos-23.11, gcc 13, CUDA 12)"
    echo "date: $(date -u +%Y-%m-%dT%H:%M:%SZ)"
    for t in python g++ gcc make nvcc; do printf '%s: %s\n' "$t" "$($t --version 2>&1 | head -1)"; done
    python - <<'PY'
import importlib.metadata as md
for p in ("torch", "numpy", "pandas", "pyarrow", "pytest", "hypothesis", "nvidia-ml-py"):
    try: print(f"{p}=={md.version(p)}")
    except Exception: print(f"{p}: not installed")
PY
  } > env.lock
  echo "wrote env.lock"
}
case "${1:-}" in
  --env) write_env; exit 0 ;;
  --restore)
    f="$CACHE/b200cnn-$(key).tgz"
    [ -f "$f" ] || { echo "no cache for source key $(key) in $CACHE"; exit 1; }
    tar xzf "$f" -C "$ROOT" && echo "restored $(tar tzf "$f" | wc -l) files from $f"; exit 0 ;;
esac
python -c "import __graft_entry__ as g; g.build()"
[ -d /root/reference ] && python baseline/build_reference.py > /dev/null 2>&1 || true
write_env
mkdir -p "$CACHE"
f="$CACHE/b200cnn-$(key).tgz"
paths=(lib/libb200cnn_cpu.so lib/libb200cnn_cuda.so env.lock)
[ -d baseline/build ] && paths+=(baseline/build)
for v in v1_serial v2_mpi_only/2.1_broadcast_all v2_mpi_only/2.2_scatter_halo v3_cuda_only v4_mpi_cuda; do
  [ -x "baseline/_ref/src/final_project/$v/template" ] && paths+=("baseline/_ref/src/final_project/$v/template")
done
tar czf "$f" "${paths[@]}"
grep -qxF "build_cache_local/" .gitignore || echo "build_cache_local/" >> .gitignore
echo "cache: $f ($(du -h "$f" | cut -f1), ${#paths[@]} entries); restore with ./create_local_cache.sh --restore"
