"""BASELINE configs[4]: knot spacing (0.02 .. 0.2 s) x spline order (4, 6) at a fixed 2 M observations — band width against the cost of
the bordered band Cholesky and of the block-Jacobi PCG on the same reduced system.  One table row per (dt, order).
usage: python profiles/tools/sweep_config5.py [dt:order ...]      default: all eight"""
import importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
pkg = importlib.import_module("rpg_vision-based_slam_b200"); scene = importlib.import_module("rpg_vision-based_slam_b200.scene")
cases = [tuple(a.split(":")) for a in sys.argv[1:]] or [(dt, o) for dt in ("0.2", "0.1", "0.05", "0.02") for o in ("4", "6")]
rows = []
for dt, order in cases:
    dt, order = float(dt), int(order)
    sc = scene.config_scene(5, dt_s=dt, order=order)
    row = {"dt_s": dt, "order": order, "knots": int(sc["knot_q"].shape[0])}
    for solver in ("cholesky", "pcg"):
        g = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
        pkg.load_scene(g, sc)
        g.setOption("bench_no_termination", 1)
        g.setOption("linear_solver", 1 if solver == "pcg" else 0)
        g.setOption("pcg_max_iters", 8000)
        nw, nt = (2, 3) if dt >= 0.1 else (1, 2)       # fine knot spacings: fewer iterations (a PCG solve takes seconds there)
        try:
            s0 = g.optimize(nw)                        # warm-up: structure, graph capture
            s = g.optimize(nt)
            n = g.counts()
            row.update(obs=n["obs"], unknowns=n["cols"], nbc=int(g.getStat("band_block_columns")), nbw=int(g.getStat("band_half_width_blocks")))
            row[solver] = {"ms_per_it": 1e3 * s["time_total_s"] / nt, "assemble_ms": 1e3 * s["time_build_s"] / nt, "solve_ms": 1e3 * s["time_solve_s"] / nt,
                           "cost_after_5": float(s["final_cost"]), "lm_iterations": nw + nt, "accepted": int(s0["num_successful_steps"] + s["num_successful_steps"])}
            if solver == "pcg":
                row[solver].update(iterations_last=int(g.getStat("pcg_iterations")), rel_residual_last=g.getStat("pcg_rel_residual"))
        except pkg.CtbaError as e:
            row[solver] = {"error": str(e)[:120]}
        g.close()
    rows.append(row)
    print(json.dumps(row), flush=True)
print("| dt [s] | order | knots | unknowns | band half-width [blocks of 32] | Cholesky solve ms | PCG solve ms | PCG iterations | rel. cost difference after the LM iterations run (5 or 3) |")
print("|---|---|---|---|---|---|---|---|---|")
for r in rows:
    c, q = r.get("cholesky", {}), r.get("pcg", {})
    rel = abs(c["cost_after_5"] - q["cost_after_5"]) / c["cost_after_5"] if "cost_after_5" in c and "cost_after_5" in q else float("nan")
    print(f"| {r['dt_s']} | {r['order']} | {r['knots']} | {r.get('unknowns', '')} | {r.get('nbw', '')} | {c.get('solve_ms', float('nan')):.2f} | "
          f"{q.get('solve_ms', float('nan')):.2f} | {q.get('iterations_last', '')} | {rel:.2e} |")
