"""GPU exploration run (development aid): timings per mode, per-kernel device time, parity statistics vs the oracle."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from tesseract_b200 import capi, scene as S  # noqa: E402
from tesseract_b200.managers import BatchChecker  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
    lib = capi.load_library()
    tf = C.c_double()
    capi.check(lib.t4c_measure_fp32_peak(0, C.byref(tf)), lib)
    print("fp32 FMA peak TFLOP/s:", tf.value)
    b = S.build_config2(n_prims=10000, margin=0.05)
    t = time.time()
    bc = BatchChecker(b)
    print("scene create s:", time.time() - t)
    q = S.random_states(b, n, seed=1234, dtype=np.float64)
    lib.t4c_profile_enable(1)
    for name, tt, want in (("first_hitonly", capi.TEST_FIRST, False), ("first_contacts", capi.TEST_FIRST, True),
                           ("closest", capi.TEST_CLOSEST, True), ("all", capi.TEST_ALL, True)):
        for rep in range(3):
            lib.t4c_profile_reset()
            t = time.time()
            hit, con, stats = bc.check_discrete(q, tt, want_contacts=want, capacity=16 * n)
            dt = time.time() - t
        prof = capi.profile_get(lib)
        print(json.dumps({"mode": name, "n": n, "wall_s": dt, "states_per_s": n / dt, "colliding": int(hit.sum()),
                          "contacts": len(con), "stats": stats,
                          "kernel_ms": {k: round(v[0], 3) for k, v in prof.items() if v[1]}}))
    # parity statistics on a prefix
    from oracle import pyoracle
    import parity
    m = min(n, 3000)
    hit_o, con_o, cnt = pyoracle.check_discrete(b, q[:m], capi.TEST_CLOSEST, n_threads=os.cpu_count() or 1)
    hit_g, con_g, _ = bc.check_discrete(q[:m], capi.TEST_CLOSEST)
    print("hit mismatches:", int((hit_o != hit_g).sum()), "contacts", len(con_g), len(con_o), "oracle counters", cnt)
    G, O = parity.index(con_g, False), parity.index(con_o, False)
    dd = [abs(G[k][0] - O[k][0]) for k in G if k in O]
    print("keys only gpu/oracle:", len(set(G) - set(O)), len(set(O) - set(G)), "max dd", max(dd), "frac>1e-5",
          float(np.mean(np.array(dd) > 1e-5)))


if __name__ == "__main__":
    main()
