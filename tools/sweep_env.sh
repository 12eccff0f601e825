#!/bin/bash
# usage: tools/sweep_env.sh <out-prefix> <bench args...> -- "ENV1=a ENV2=b" "ENV3=c" ...   (one bench run per env set)
prefix=$1; shift
args=()
while [ "$1" != "--" ]; do args+=("$1"); shift; done
shift
i=0
for envset in "$@"; do
  env $envset python bench.py "${args[@]}" --no-cpu-baseline > "${prefix}_$i.log" 2>&1
  python - "$envset" "${prefix}_$i.log" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    k = d["roofline"]["kernel_ms_per_step"]
    print(sys.argv[1] or "(default)", "| value %.2fM e2e %.2fM |" % (d["value"] / 1e6, d["e2e"]["value"] / 1e6),
          " ".join("%s %.2f" % (n, v) for n, v in k.items()))
except Exception as e:
    print(sys.argv[1], "ERR", e)
PY
  i=$((i+1))
done
