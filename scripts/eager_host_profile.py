"""Tuning aid: cProfile of the host side of one eager (no CUDA graph) search at the headline shape - where the Python time of
the driver loop goes between kernel launches."""
import cProfile, pstats, sys, os, time
sys.path.insert(0, os.getcwd())
import bench, torch
sys.argv=["bench.py","--no-graph"]
a=bench.parse_args(); dev=torch.device("cuda",0)
inp=bench.make_inputs(a,0,dev); model=bench.TableModel(inp,a.variant,a.actions); mcts=bench.make_driver(a,dev,False)
for _ in range(3): bench.one_search(a,mcts,model,inp)
torch.cuda.synchronize()
t0=time.perf_counter()
for _ in range(10): bench.one_search(a,mcts,model,inp)
t1=time.perf_counter(); torch.cuda.synchronize()
print("host enqueue ms/search %.3f"%((t1-t0)/10*1e3))
pr=cProfile.Profile(); pr.enable()
for _ in range(10): bench.one_search(a,mcts,model,inp)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
