import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["DL_ESM_ALIGNMENT"] = "2"
import dl_esm_inf_b200 as fd
fd.gocean_initialise(1, 1, device=0)
bcs = (fd.GO_BC_PERIODIC, fd.GO_BC_PERIODIC, fd.GO_BC_NONE)
g = fd.grid_type(fd.GO_ARAKAWA_C, bcs, fd.GO_OFFSET_SW)
g.decompose(256, 64)
fd.grid_init(g, 1e5, 1e5)
sw = fd.shallow_water(g)
sw.set_variant(int(sys.argv[1]) if len(sys.argv) > 1 else 2)
sw.init_device()
sw.step(2)
print("checksums", sw.checksums())
