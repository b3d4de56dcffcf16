timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for f in 0 1 5; do echo "== SFC_ACL_FEAT=$f"; SFC_ACL_FEAT=$f timeout 600 python tools/bench_configs.py cfg4 2>&1 | head -1 | cut -c1-260; done
echo "== auto"; timeout 600 python tools/bench_configs.py cfg4 2>&1 | head -1 | cut -c1-260
