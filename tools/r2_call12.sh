# 1-GPU call: what bounds the dense transform -- ring depth (2 / 3 / 4 slots), tile width 64 (6 slots), drain interval
mkdir -p gpurun_out
for v in "gemm_stages=2" "gemm_stages=3" "gemm_stages=4" "gemm_bn=64"; do
  timeout 200 python tools/gemm_bench.py --tune $v > gpurun_out/r2l_gemm_$v.jsonl 2>&1; echo "$v rc=$?"
done
for k in 4 8 32; do
  timeout 200 python tools/gemm_bench.py --kgroup $k > gpurun_out/r2l_gemm_kgroup=$k.jsonl 2>&1; echo "kgroup=$k rc=$?"
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r2l_gemm_*.jsonl")):
    tot = 0; rows = []
    for l in open(f):
        try: d = json.loads(l)
        except Exception: print(f, l[:300]); continue
        if d["tag"].startswith("tcgen05"):
            tot += d["ms"]
            rows.append("%d/%d/%d:%.3f(%.1e)" % (d["M"] if d["M"] < 9999 else 0, d["N"], d["K"] if d["K"] < 9999 else 0, d["ms"], d["max_rel_err"]))
    print(f.split("/")[-1], "sum %.3f" % tot, " ".join(rows))
PY
