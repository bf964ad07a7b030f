import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg = g.load_package()
with pkg.Context() as c:
    for _ in range(3):
        d = c.measure_int_peaks()
        mhz = d["sm_mhz"]
        print({k: (round(v / 148 / (mhz * 1e6), 2) if k != "sm_mhz" else v) for k, v in d.items()})
