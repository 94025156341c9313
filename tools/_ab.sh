python -m pytest tests/test_gpu_volume.py -x -q 2>&1 | tail -2
for i in 1 2 3; do
  echo prev; PP_B200_LIB=path_planner_b200/libpp_b200_prev.so python tools/k1_probe.py rect 20
  echo new; python tools/k1_probe.py rect 20
done
