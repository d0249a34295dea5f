"""Does a CUDA graph of the acoustic sub-step pay for the small reference configs?  The device-resident loop
(dt / relaxation time in a device control block) makes the sub-step's launches identical from one sub-step to the next,
so one sub-step is captured (torch.cuda.CUDAGraph on the stream the library works on) and replayed.
    python scratch/graph_experiment.py > gpurun_out/graph_experiment.json"""
import importlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
pk = importlib.import_module("sphinxsys-learning_b200")
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
out = {}
for name, ndim, dp, order in (("C1 2-D dam break, 3200 fluid", 2, 0.025, 0), ("C3 3-D dam break default, 8000 fluid", 3, 0.05, 1),
                              ("3-D dam break, 1.0 M fluid", 3, 0.01, 1)):
    c = pk.cases.DamBreak(pk.host, ndim=ndim, dp=dp, with_observer=False, device=0, stream=stream.cuda_stream)
    S, w = c.sys, c.water
    for _ in range(10):
        c.advection_step()
    def substep():
        S.integration_1st_half(c.inner, c.wall_contact, True, -1.0)
        S.integration_2nd_half(c.inner, c.wall_contact, True, -1.0)
        S.substep_ctl(w, 0.6, order, 1)
    S.relaxation_begin(1.0e9, c.dt)       # Dt far away: the loop never finishes, every replay is a live sub-step
    for _ in range(5):
        substep()
    torch.cuda.synchronize()
    N = 2000 if w.n < 100000 else 200
    l0 = S.launches
    t0 = time.perf_counter()
    for _ in range(N):
        substep()
    torch.cuda.synchronize()
    eager = (time.perf_counter() - t0) / N * 1e6
    per = (S.launches - l0) / N
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        substep()
    torch.cuda.synchronize()
    for _ in range(5):
        g.replay()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(N):
        g.replay()
    torch.cuda.synchronize()
    graph = (time.perf_counter() - t0) / N * 1e6
    g4 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g4, stream=stream):
        for _ in range(4):
            substep()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(N // 4):
        g4.replay()
    torch.cuda.synchronize()
    graph4 = (time.perf_counter() - t0) / (N // 4 * 4) * 1e6
    st = S.relaxation_state()
    out[name] = {"n_fluid": w.n, "launches_per_substep": per, "us_per_substep_eager": round(eager, 1), "us_per_substep_graph": round(graph, 1),
                 "us_per_substep_graph_of_4": round(graph4, 1), "substeps_counted_by_the_device": st["substeps"]}
    print(json.dumps({name: out[name]}), file=sys.stderr, flush=True)
print(json.dumps(out, indent=1))
