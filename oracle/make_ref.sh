#!/bin/sh
# Recipe for oracle/_ref/: the reference's OWN native aligner, as shipped.
#
# GRIDSS ships its Smith-Waterman arithmetic only as a prebuilt binary
# (/root/reference/src/main/resources/libsswjni.so; no C sources in the tree),
# so "building the reference" here means staging that binary, unmodified, next to a small C driver
# (oracle/ref_driver.c, ours) that calls it with exactly the arguments the Java wrapper passes.
# oracle/_ref/ is git-ignored (stays out of history) but travels to the GPU box with gpurun.
# TEST/BENCH INFRASTRUCTURE ONLY - nothing under gridss_b200/ may load anything from oracle/.
set -e
here="$(cd "$(dirname "$0")" && pwd)"
mkdir -p "$here/_ref"
src=/root/reference/src/main/resources/libsswjni.so
if [ -f "$src" ]; then
    cp "$src" "$here/_ref/libsswjni.so"
    sha1sum "$here/_ref/libsswjni.so" | cut -d' ' -f1 > "$here/_ref/libsswjni.sha1"
fi
# ours: the dlopen driver and the scalar restatement
gcc -O2 -fPIC -shared -pthread -o "$here/_ref/libref_driver.so" "$here/ref_driver.c" -ldl
gcc -O3 -march=x86-64-v2 -fPIC -shared -pthread -o "$here/libssw_model.so" "$here/ssw_model.c"
echo "oracle built: $(ls "$here/_ref")"
