export SGM_B200_LIBRARY=$PWD/stereo_matching_b200/libsgm_b200_tuning.so
python tests/run_one.py sgm 375 1242 128 3 2>&1 | tail -n 4
python tests/run_one.py bmsad 375 1242 64 3 2>&1 | tail -n 3
python tests/run_one.py sgm 1080 1920 256 3 2>&1 | tail -n 3
