# shade-kernel occupancy variants (alternative builds under gpurun_variants/): identical films, Msamples/s
for lib in "" $(ls gpurun_variants/*.so 2>/dev/null); do
  for sc in "scenes/classroom/vision_scene.json 0 16" "scenes/cbox/cbox_dispersion.json 1024 64" "scenes/cbox/cbox_principled.json 1024 32"; do
    VISION_B200_LIB=$lib python tools/sweep_shade.py $sc 2>&1 | tail -1
  done
done
