ALLOFF="MEEMB200_NO_STREAM=1 MEEMB200_NO_STREAM_SYRK=1 MEEMB200_NO_BATCH_INV=1"
for nm in 100 150 90 200; do echo "== modes $nm (vs all-off baseline)"; timeout 120 python tools/experiments/stream_check.py 512 $nm $ALLOFF 2>&1 | tail -7 | cut -c1-200; done
