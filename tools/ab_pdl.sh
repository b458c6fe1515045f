# A/B of programmatic dependent launch for the step kernels on one box, alternating runs.
# SWARM_PDL: 0 = plain launches, 1 = trigger at the kernel's top, late = trigger after its main loop, auto = the library's choice.
B="python bench.py --steps 256 --warmup 5 --no-cpu-baseline --no-strong --no-rollout --e2e-steps 2"
for rep in 1 2; do
  for cfg in "cfg3" "cfg3 --envs 8192" "cfg3 --envs 32768" "cfg5b" "cfg2" "cfg4" "cfg5a"; do
    tag=$(echo $cfg | tr -d ' -')
    for mode in 0 1 late auto; do
      if [ $mode = auto ]; then unset SWARM_PDL; else export SWARM_PDL=$mode; fi
      $B --workload $cfg > gpurun_out/ab_${mode}_${tag}_$rep.json 2>> gpurun_out/ab_err.log || echo FAIL $mode $cfg
    done
  done
done
python - <<'PY'
import json, glob, collections
r = collections.defaultdict(list)
for f in sorted(glob.glob("gpurun_out/ab_*_*_?.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as ex:
        print(f, "ERR", ex); continue
    _, mode, tag, rep = f.split("/")[-1][:-5].split("_")
    r[(tag, mode)].append(round(d["ms_per_step"] * 1e3, 3))
for k in sorted(r): print(k, r[k])
PY
