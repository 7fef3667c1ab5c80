# Sustained stream (4096 pre-staged frames) + run-to-run spread of the default bench line.
mkdir -p gpurun_out
python bench.py --steps 93 --warmup 5 --no-cpu 2>&1 | tail -1 > gpurun_out/stream_4092.json
for i in 1 2 3 4 5; do python bench.py --no-cpu 2>&1 | tail -1 > gpurun_out/repeat_$i.json; done
python - <<'PY'
import json
r = json.load(open("gpurun_out/stream_4092.json"))
print("stream: %d frames, value %.0f, e2e %.0f, clocks %s" % (r["steps"] * 44, r["value"], r["e2e"]["value"], r["clocks"]))
v = sorted(json.load(open("gpurun_out/repeat_%d.json" % i))["value"] for i in range(1, 6))
e = sorted(json.load(open("gpurun_out/repeat_%d.json" % i))["e2e"]["value"] for i in range(1, 6))
print("5 runs: value min/median/max %.0f / %.0f / %.0f, e2e %.0f / %.0f / %.0f" % (v[0], v[2], v[4], e[0], e[2], e[4]))
PY
