python tools/c3_tune.py 4 4
B200_FFT_COLS=25 python tools/c3_tune.py 4 4
B200_FFT_COLS=20 python tools/c3_tune.py 4 4
python tools/acq_tune.py C2 50
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
