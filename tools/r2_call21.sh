# 1-GPU call: GEMM (TS form) with the truncating hi/lo split against the rounding one
mkdir -p gpurun_out
timeout 150 python tools/gemm_bench.py --split 0 1 --kgroup 2 4 > gpurun_out/r2u_gemm_split.jsonl 2>&1; echo "rc=$?"
python - <<'PY'
import json
tot = {}
for l in open("gpurun_out/r2u_gemm_split.jsonl"):
    try: d = json.loads(l)
    except Exception: print(l[:200]); continue
    if d["tag"].startswith("tcgen05"):
        k = (d["split"], d["kgroup"]); tot.setdefault(k, [0, 0]); tot[k][0] += d["ms"]; tot[k][1] = max(tot[k][1], d["max_rel_err"])
        print("%s/%s/%s split %d kg %d: %.4f ms %.1e" % (d["M"] if d["M"] < 9999 else "n", d["N"], d["K"] if d["K"] < 9999 else "n", d["split"], d["kgroup"], d["ms"], d["max_rel_err"]))
print({k: (round(v[0], 4), v[1]) for k, v in tot.items()})
PY
