# BVH pre-splitting sweep (run under gpurun): nodes / triangles per ray and stage times of a 2-spp classroom / dragon render
for scene in "scenes/classroom/vision_scene.json 1920 1080" "scenes/cbox_dragon/vision_scene.json 1024 1024"; do
  for ploc in -1 0 1; do
    for split in 0 15 30 60 120; do
      echo "=== $scene ploc=$ploc split=$split"
      VB_COUNT=1 VB_PLOC=$ploc VB_SPLIT=$split python tools/prof_pt.py $scene 2 2 2>&1 | grep -E "build:|work:|rep 1|trace_closest" | tail -4
    done
  done
done
