import sys, json
sys.argv = ["bench.py", "--no-cpu-baseline", "--no-stream", "--steps", "20"]
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
import bench
bench.CFG["ransac_max_iters"] = int(sys.stdin.readline() or 2000)
import io, contextlib
buf = io.StringIO()
with contextlib.redirect_stdout(buf):
    bench.main()
d = json.loads(buf.getvalue().strip().splitlines()[-1])
print("iters", bench.CFG["ransac_max_iters"], "value", round(d["value"], 1), "ransac", round(d["stages"]["ransac"]["ms_per_step"], 3), "pose", round(d["stages"]["pose"]["ms_per_step"], 3))
