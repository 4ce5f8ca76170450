import sys, os, time, ctypes as C
sys.path.insert(0, os.getcwd())
import bench
from easyav1_b200.lib import load_host
man, obu = bench.load_clip()
host = load_host()
cap = man["w"] * man["h"] * 4
buf = C.create_string_buffer(cap)
def run(tag):
    for _ in range(2): bench.e2e_pass(host, obu, cap, buf)
    t = 0; n = 0
    for _ in range(4):
        nn, dt, md5, cnt = bench.e2e_pass(host, obu, cap, buf); t += dt; n += nn
    print(tag, "%.1f fps" % (n / t))
run("no torch:")
import torch
torch.cuda.set_device(0)
x = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
run("torch imported + cuda init:")
torch.set_num_threads(1)
run("torch 1 thread:")
