# bench the default library and every development variant given as argument (lib suffixes)
for v in "" "$@"; do
  PSFM_B200_LIB=$PWD/packnet-sfm-resnet-san_b200/lib/libpsfm_b200$v.so python bench.py --steps 100 --warmup 10 --no-cpu 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('variant [$v]', round(d['value']), 'frames/s', round(d['ms_per_step'],4), 'ms/step', d['roofline']['all_kernels_ms_per_step'])"
done
