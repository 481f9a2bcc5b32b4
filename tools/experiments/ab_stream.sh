ALLOFF="MEEMB200_NO_STREAM=1 MEEMB200_NO_STREAM_SYRK=1 MEEMB200_NO_BATCH_INV=1"
for nm in 200 128 104 72 248 256 88 40; do echo "== modes $nm (vs all-off baseline)"; timeout 120 python tools/experiments/stream_check.py 512 $nm $ALLOFF 2>&1 | tail -7 | cut -c1-200; done
echo "== full: all on vs all off"; timeout 200 python tools/experiments/stream_check.py 4096 200 $ALLOFF 2>&1 | tail -7
echo "== full: all on vs no stream gemm1"; timeout 200 python tools/experiments/stream_check.py 4096 200 MEEMB200_NO_STREAM=1 2>&1 | tail -7 | head -2
echo "== full: all on vs no stream syrk"; timeout 200 python tools/experiments/stream_check.py 4096 200 MEEMB200_NO_STREAM_SYRK=1 2>&1 | tail -7 | head -2
echo "== full: all on vs syrk kc8"; timeout 200 python tools/experiments/stream_check.py 4096 200 MEEMB200_SYRK_KC8=1 2>&1 | tail -7 | head -2
