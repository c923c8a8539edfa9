# compare alternative builds of libvision_b200 (gpurun_variants/*.so) on classroom, dispersion and the microbench
for lib in "" $(ls gpurun_variants/*.so 2>/dev/null); do
  echo "== ${lib:-default}"
  VISION_B200_LIB=$lib python tools/sweep_shade.py scenes/classroom/vision_scene.json 0 16 2>&1 | tail -1 | cut -c1-260
  VISION_B200_LIB=$lib python tools/prof_trace.py 10000000 16777216 4 2>&1 | tail -1
done
