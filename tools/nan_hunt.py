"""Find ships that leave the physical regime in the bench workload and replay one on the CPU oracle."""
import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np, torch
import bench, helpers as H
import sac_maritime_ast_b200 as sit
from oracle import cport

S, KI, MAXWP, NL = 1 << 16, 256, 8, 12
rng = np.random.default_rng(bench.SEED)
n_wp, wp_n, wp_e, psi0, u_sp = bench.make_routes(S, MAXWP, rng)
dists = [bench.make_dist(S, rng) for _ in range(NL)]
dev = torch.device("cuda:0")
sm, tc, ap = sit.default_ship([(0, 0), (1000, 1000)], 0.0)
for prec in ("fp32", "fp64"):
    fl = sit.ShipFleet(sm, tc, ap, 8.0, n_ships=S, device=dev, precision=prec, max_wp=MAXWP)
    fl.set_routes(torch.from_numpy(wp_n).to(dev), torch.from_numpy(wp_e).to(dev), torch.from_numpy(n_wp).to(dev))
    fl.f["north"].copy_(torch.from_numpy(wp_n[0].astype(np.float64))); fl.f["east"].copy_(torch.from_numpy(wp_e[0].astype(np.float64)))
    fl.f["psi"].copy_(torch.from_numpy(psi0)); fl.f["u_sp"].copy_(torch.from_numpy(u_sp))
    first_bad = None
    for j in range(NL):
        fl.step(KI, dist=torch.from_numpy(dists[j]).to(dev))
        u = fl.f["u"]; om = fl.f["omega"]
        bad = ~torch.isfinite(u) | (u.abs() > 50) | (om < 0)
        print(prec, "launch", j, "bad ships", int(bad.sum()), "nan", int((~torch.isfinite(u)).sum()), "omega<0", int((om < 0).sum()),
              "u range", float(u[torch.isfinite(u)].min()), float(u[torch.isfinite(u)].max()))
        if first_bad is None and bad.any():
            first_bad = (j, int(torch.nonzero(bad)[0]))
    print(prec, "first bad", first_bad)

if first_bad is not None:
    j_bad, i = first_bad
    cfg = H.default_cfg(); p = cport.make_params(cfg); L = cport.lib()
    s = cport.OrcShip()
    route = [(float(wp_n[j, i]), float(wp_e[j, i])) for j in range(n_wp[i])]
    cport.init_ship(s, cfg, route, float(psi0[i]), float(u_sp[i]))
    rud, thr = C.c_double(), C.c_double()
    print("ship", i, "u_sp", u_sp[i], "n_wp", n_wp[i], "route", route[:3])
    for j in range(j_bad + 1):
        s.vc_n, s.vc_e, s.wind_speed, s.wind_dir = [float(x) for x in dists[j][:, i]]
        for t in range(KI):
            L.orc_ship_step(C.byref(p), C.byref(s), 0, 1, None, C.byref(rud), C.byref(thr))
            if (j == j_bad and t % 16 == 0) or not np.isfinite(s.u) or s.omega < 1.0:
                print(j, t, "u %.4g v %.4g om %.4g thr %.4g eis %.4g eiw %.4g k %d dist %s" % (s.u, s.v, s.omega, thr.value, s.ei_speed, s.ei_shaft, s.k, dists[j][:, i]))
            if not np.isfinite(s.u):
                raise SystemExit
