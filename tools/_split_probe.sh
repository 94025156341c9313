python -m pytest tests/test_gpu_volume.py -x -q 2>&1 | tail -3
BANDS=512,1024,2048,4096 python tools/band_probe.py 10
python tools/k1_probe.py box 10
python tools/k1_probe.py solid 10
