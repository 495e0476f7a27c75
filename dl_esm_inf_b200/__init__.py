"""dl_esm_inf_b200 -- B200-native drop-in for the hot path of stfc/dl_esm_inf (fd_lib /
dm_fd_lib): stencil sweeps over r2d_field data on an Arakawa C-grid grid_type, and the halo
exchange between them.  The product is libdlesm_b200.so (hand-written sm_100a CUDA + C++
host code behind the C ABI of include/dlesm_b200.h); this package is the thin host-side
mirror of the reference's Fortran API used where no Fortran toolchain exists."""
from .api import *  # noqa: F401,F403
from .api import (GoceanStop, Comms, Region, Subdomain, comm_unique_id, copy_field, decomposition_type,  # noqa: F401
                  dev_sync, field_checksum, free_field, get_num_ranks, get_rank, global_sum,
                  go_decompose, gocean_finalise, gocean_initialise, gocean_stop, grid_init,
                  grid_type, halo_exchange_fields, iprocmap, kernel_launch_count, make_region,
                  map_comms, on_master, r2d_field, set_extensions, set_field, shallow_water)

__version__ = "0.1.0"
