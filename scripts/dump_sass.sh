#!/bin/bash
# SASS listings of the shipped kernels (cuobjdump on the built library) -> profiles/sass/
set -e
cd "$(dirname "$0")/.."
mkdir -p profiles/sass
cuobjdump -sass libgolb200.so > /tmp/golb200_all.sass
split_fn() {  # $1 = regex on the mangled "Function :" line, $2 = output file (encodings stripped)
  awk -v pat="$1" '/Function : /{f = ($0 ~ pat)} f' /tmp/golb200_all.sass > "profiles/sass/$2"
  echo "$2: $(grep -c . profiles/sass/$2) lines, $(grep -cE '^\s+/\*[0-9a-f]{4}\*/' profiles/sass/$2) instructions"
}
split_fn 'life_strip_kernelILi8ELb0E' life_strip_k8.sass
split_fn 'life_strip_kernelILi8ELb1E' life_strip_k8_flags.sass
split_fn 'halo_push_kernel|scan_alive_kernel|gather_flags_kernel|dense_to_padded|padded_to_dense|popcount_padded|hash_padded' engine_aux_kernels.sass
split_fn 'pack_kernel|unpack_kernel|step_dense_kernel|flags_dense_kernel|random_dense_kernel|popcount_kernel' dense_kernels.sass
cuobjdump -sass bin/gol-tilebench 2>/dev/null | awk '/Function : /{f = ($0 ~ /life_tile_kernelILi8ELi34ELi8ELi1ELi1ELi2E|life_tile_kernelILi8ELi34ELi8ELi2ELi2ELi2E/)} f' > profiles/sass/experiment_tile_tma_cluster.sass
echo "experiment_tile_tma_cluster.sass: $(grep -c . profiles/sass/experiment_tile_tma_cluster.sass) lines; TMA / cluster instructions:"
grep -oE "UTMALDG[.A-Z0-9]*|UBLKCP[.A-Z0-9]*|SYNCS[.A-Z0-9]*|UCGABAR_[A-Z]*|LDS[.A-Z0-9]*" profiles/sass/experiment_tile_tma_cluster.sass | sort | uniq -c
# keep the instruction text only (drop the encoding words: a third of the size)
python - <<'PY'
import glob, re
for f in glob.glob("profiles/sass/*.sass"):
    out = []
    for l in open(f):
        if re.match(r"^\s+/\* 0x[0-9a-f]+ \*/\s*$", l):
            continue
        l = re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l.rstrip("\n"))
        l = re.sub(r"^\s+(/\*[0-9a-f]+\*/)\s+", r"  \1 ", l)
        if l.strip():
            out.append(l)
    open(f, "w").write("\n".join(out) + "\n")
PY
