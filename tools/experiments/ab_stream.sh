ALLOFF="MEEMB200_NO_STREAM=1 MEEMB200_NO_STREAM_SYRK=1 MEEMB200_NO_BATCH_INV=1"
for nm in 200 104 128 72 248; do echo "== modes $nm (vs all-off baseline)"; timeout 120 python tools/experiments/stream_check.py 512 $nm $ALLOFF 2>&1 | tail -7 | cut -c1-200; done
echo "== full: default vs all off"; timeout 200 python tools/experiments/stream_check.py 4096 200 $ALLOFF 2>&1 | tail -7
