for L in "" build_variants/librt_dblshf.so; do
env ${L:+RT_B200_LIB=$PWD/$L} timeout 200 python tools/run_once.py --spp 500 --reps 3 2>&1 | tail -1 | python -c "import sys,json; [print(round(json.loads(l)['render_ms'],2), json.loads(l)['image_sha256'], json.loads(l)['rays']) for l in sys.stdin]"
done
