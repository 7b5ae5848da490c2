#!/bin/bash
# rollout experiments: prints value, ms per step and the stage table of one short C2 bench run per setting.
#   tools/parts_sweep.sh [parts ...]     TSM_COLLECT_STREAMS = number of world ranges the lock-step rollout runs on separate streams
#   other switches (TSM_COLLECT_NO_PDL=1, TSM_B200_LIB=...) are inherited from the environment
for p in ${@:-2}; do
  TSM_COLLECT_STREAMS=$p python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-c5 2>/dev/null > /tmp/parts_$p.json
  python - "$p" <<'PY'
import json, sys
p = sys.argv[1]
d = json.loads(open(f"/tmp/parts_{p}.json").read())
st = d.get("stages_ms") or {}
print("parts", p, round(d["value"] / 1e6, 1), "M", round(d["ms_per_step"], 2), "ms", {k: round(v["ms_per_step"], 2) for k, v in st.items()})
PY
done
