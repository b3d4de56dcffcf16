export SFC_FLAGS=2
for cfg in "128x2 2 -1" "128x2 1 -1" "256x2 1 -1" "256x3 2 -1" "128x3 3 -1" "256x4 2 -1"; do
  set -- $cfg
  echo "== variant $1 ctas $2 carveout $3"
  SFC_VARIANTS=$1 SFC_CTAS=$2 SFC_CARVEOUT=$3 timeout 200 python tools/tune.py empty cfg2 cfg3u 2>&1 | grep -v "^#"
done
echo "== serial launches (no overlap), 128x2 with 3 / 9 ctas"
SFC_FLAGS=0 SFC_VARIANTS=128x2 SFC_CTAS=3 timeout 200 python tools/tune.py cfg2 2>&1 | grep -v "^#"
SFC_FLAGS=0 SFC_VARIANTS=128x2 SFC_CTAS=0 timeout 200 python tools/tune.py cfg2 2>&1 | grep -v "^#"
