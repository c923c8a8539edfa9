for lib in "" gpurun_variants/libvb_mb5.so gpurun_variants/libvb_mb6.so; do
  for sc in "scenes/classroom/vision_scene.json 0 16" "scenes/cbox/cbox_dispersion.json 1024 64"; do
    VISION_B200_LIB=$lib python tools/sweep_shade.py $sc 2>&1 | tail -1
  done
done
