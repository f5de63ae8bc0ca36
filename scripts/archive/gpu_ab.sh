# A/B of kernel switches on one box: AB_SHAPE="--worlds .." ; AB_ENVS="A=1,B=2 C=3 ..." ('+'-separated env per variant, '-' = none)
for v in ${AB_ENVS}; do
  envs=$(echo $v | tr '+' ' '); [ "$v" = "-" ] && envs=""
  env $envs python bench.py --steps ${STEPS:-40} --warmup 5 --no-cpu-baseline --no-e2e $AB_SHAPE 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print('$v', '| kernel_ms', r['kernel_ms'], 'GB/s', r['achieved'], 'frac', r['frac'])"
done
