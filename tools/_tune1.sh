python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python tools/acq_tune.py C1 32
python tools/acq_tune.py C4_gps 32
python tools/acq_tune.py C4_gal 36
python tools/acq_tune.py C4_bds 63
python tools/acq_tune.py C2 50
B200_ACQ_FWD_GENERIC=1 python tools/acq_tune.py C2 50
python tools/c3_tune.py 4 4
