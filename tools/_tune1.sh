for nb in 5 10 20 40 80; do B200_FFT_NB=$nb python tools/acq_tune.py C2 50; done
for nb in 5 10 20 40 80; do B200_FFT_NB=$nb python tools/acq_tune.py C2 50; done
for nb in 4 5 10 20 25; do B200_FFT_NB=$nb python tools/acq_tune.py C4_gal 36; done
for nb in 5 10 20 25 50; do B200_FFT_NB=$nb python tools/acq_tune.py C4_bds 63; done
