import json, os, sys, time
sys.path.insert(0, os.getcwd())
import __graft_entry__ as ge
gp = ge.load_package().gpuprobe
n = 1 << 28
a, oa = gp.pinned_array(n); b, ob = gp.pinned_array(n); c, oc_ = gp.pinned_array(n)
a[:] = 1.0; b[:] = 2.0
def run(tag):
    c[:] = 0
    gp.vectoradd_buffers(a, b, c, n)
    t0 = time.perf_counter()
    for _ in range(5): gp.vectoradd_buffers(a, b, c, n)
    dt = (time.perf_counter() - t0) / 5
    assert c[0] == 3.0 and c[-1] == 3.0 and c[n//2] == 3.0
    print(tag, round(dt*1e3, 2), "ms", round(12*n/dt/1e9, 2), "GB/s", flush=True)
run("staged")
os.environ["GPUPROBE_BUF_ZEROCOPY"] = "1"
for v in (11, 1, 2, 3, 12, 5, 7):
    os.environ["GPUPROBE_BUF_ZEROCOPY_VARIANT"] = str(v)
    run("zerocopy variant %s" % gp.VARIANT_NAMES[v])
