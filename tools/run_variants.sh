for v in base probe2; do
  if [ $v = base ]; then unset B200RL_LIB; else export B200RL_LIB=/root/repo/tools/var_$v.so; fi
  timeout 200 python bench.py --steps 10 --warmup 3 --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', d['phases'])"
done
