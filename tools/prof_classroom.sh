set -x
python tools/prof_pt.py scenes/classroom/vision_scene.json 1920 1080 2 2 > gpurun_out/r2_prof_classroom_plain.log 2>&1 || exit 1
VB_COUNT=1 python tools/prof_pt.py scenes/classroom/vision_scene.json 1920 1080 1 1 >> gpurun_out/r2_prof_classroom_plain.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_wf_shade --launch-skip 1 --launch-count 2 -o gpurun_out/r2_classroom_shade -f python tools/prof_pt.py scenes/classroom/vision_scene.json 1920 1080 1 1 > gpurun_out/ncu_a.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_trace --launch-skip 2 --launch-count 2 -o gpurun_out/r2_classroom_trace -f python tools/prof_pt.py scenes/classroom/vision_scene.json 1920 1080 1 1 > gpurun_out/ncu_b.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_classroom.csv python tools/prof_pt.py scenes/classroom/vision_scene.json 1920 1080 1 1 > gpurun_out/ncu_c.log 2>&1
cat gpurun_out/r2_prof_classroom_plain.log
