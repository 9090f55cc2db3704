import sys; sys.path.insert(0,'/root/repo')
from otedama_b200 import miner
for i,n in enumerate(miner.int_peak_modes()):
    if i < 9 and '--all' not in sys.argv: continue
    ops,mhz,ms = max(miner.int_peak(i, iters=4096) for _ in range(3))
    print(f"{n:14s} {ops/1e12:7.3f} Tops/s {mhz:7.1f} MHz {ops/(mhz*1e6)/148:6.1f} /clk/SM")
