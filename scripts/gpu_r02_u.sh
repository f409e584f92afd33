#!/bin/bash
# per-launch breakdown of the UFM-Base step (events around every launch, three profiled passes)
mkdir -p gpurun_out
rm -f gpurun_out/dump_base16.csv
UFM_PROFILE_DUMP=gpurun_out/dump_base16.csv timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-torch-gpu > gpurun_out/bench_u.json 2>gpurun_out/bench_u.err
echo "== exit $?"; wc -l gpurun_out/dump_base16.csv
python - <<'PY'
import collections, re
agg = collections.OrderedDict()
n = 0
for line in open("gpurun_out/dump_base16.csv"):
    parts = line.rstrip("\n").rsplit(",", 4)
    lab, cat, ms, fl, by = parts[0], int(parts[1]), float(parts[2]), float(parts[3]), float(parts[4])
    key = re.sub(r"\d+", "#", lab)
    a = agg.setdefault(key, [0, 0.0, 0.0, 0.0])
    a[0] += 1; a[1] += ms; a[2] += fl; a[3] += by
    n += 1
passes = 3
tot = sum(a[1] for a in agg.values()) / passes
print("total profiled ms per pass", round(tot, 2))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    ms = a[1] / passes
    print(f"{k:60s} n={a[0] // passes:4d} ms={ms:7.3f} share={ms / tot:6.3f} TFLOP/s={a[2] / a[1] / 1e9 if a[1] else 0:8.1f} GB/s={a[3] / a[1] / 1e6 if a[1] else 0:8.1f}")
PY
# N = 128 convs as CTA pairs (pair mode bit 1) against single-CTA tiles, final binary
timeout 600 python scripts/gemm_bench.py --only proc --variants 1:1:1:0:0:1,1:3:1:0:0:1,1:1:1:0:4:1,1:3:1:0:4:1 > gpurun_out/gemm_bench_u.txt 2>&1
cat gpurun_out/gemm_bench_u.txt
