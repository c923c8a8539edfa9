for t in test_soup_as_one_instance_and_device_rays test_many_instances_of_few_meshes_vs_oracle_and_flattened test_path_tracer_films_identical test_classroom_through_the_host_layer; do
  echo "== $t"; timeout 100 python -m pytest tests/test_two_level_gpu.py -x -q -k $t 2>&1 | tail -12; echo "rc=$?"
done
