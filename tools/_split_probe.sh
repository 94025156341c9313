for t in 8 9 10 11 12 13; do echo "tables $t"; PP_SE2_MAX_TABLES=$t BANDS=512 python tools/band_probe.py 20; done
echo no tma; PP_SE2_NO_TMA=1 BANDS=512 python tools/band_probe.py 20
echo rows32; PP_SE2_TILE_ROWS=32 BANDS=512 python tools/band_probe.py 20
