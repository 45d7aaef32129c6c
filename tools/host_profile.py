"""cProfile of one DFQ pipeline pass (host-side overhead hunting).  python tools/host_profile.py"""
import contextlib, cProfile, io, os, pstats, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
case = bench.PipelineCase(dev)
with contextlib.redirect_stdout(io.StringIO()):
    for _ in range(3):
        case.rearm(False); case.step()
    torch.cuda.synchronize()
    case.rearm(False)
    pr = cProfile.Profile()
    pr.enable()
    case.step()
    torch.cuda.synchronize()
    pr.disable()
st = pstats.Stats(pr, stream=sys.stdout)
st.sort_stats(sys.argv[1] if len(sys.argv) > 1 else "cumulative").print_stats(40)
