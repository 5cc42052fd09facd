import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools")); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np, synth
from b200_import import load_package
from b200_import import load_oracle; om = load_oracle()
from test_oracle_trk import pilot_cfg, pilot_codes
b200 = load_package(); O = om.Oracle("own")
kind, ext, n_ms = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
fs = 15e6 if kind == "gps_l5" else 4e6
p, d, dl = 7, 1230.0, 333.4
iq, t, carrier = synth.tracking_stream_pilot_data(kind, p, fs, n_ms, seed=37, cn0=50.0, doppler=d, delay=dl)
o = pilot_cfg(om.TrkCfg, kind, fs, carrier, ext)
g = b200.TrkCfg()
for name, _ in o._fields_: setattr(g, name, getattr(o, name))
S0 = int(1.2 * fs); T = fs * g.code_period_s; n_ep = int(n_ms / (1e3 * g.code_period_s)) - 3
trk = b200.TrackingLoopB200(g); pilot, data = pilot_codes(kind, p)
trk.set_code(0, pilot); trk.set_data_code(0, data); trk.start(0, float(np.fmod(S0 + dl + 0.15, T)), d - 10.0, 0, S0 + 2000)
got = trk.run(iq, S0, np.arange(1), n_ep)[:, 0]
ot = O.trk(o); ot.set_code(pilot); ot.set_data_code(data); ot.start(float(np.fmod(S0 + dl + 0.15, T)), d - 10.0, 0, S0 + 2000)
want, done = ot.run(iq, S0, n_ep)
for e in range(0, n_ep, max(1, n_ep // 60)):
    a, b = got[e], want[e]
    print(e, int(b["state"]), "dcorr=%.4f daccu=%.4f ddop=%.2e dcf=%.2e dcerr=%.2e dperr=%.2e | P=%.0f" % (np.abs(a["corr"] - b["corr"]).max(), np.abs(a["accu"] - b["accu"]).max(),
          a["carrier_doppler_hz"] - b["carrier_doppler_hz"], a["code_freq_chips"] - b["code_freq_chips"], a["code_error_chips"] - b["code_error_chips"], a["carr_error_hz"] - b["carr_error_hz"], np.hypot(b["corr"][4], b["corr"][5])))
