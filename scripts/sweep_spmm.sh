#!/bin/bash
# usage: scripts/sweep_spmm.sh  -- tile-shape sweep of k_spmm_tiled on the C4 scene (prints GB/s per config)
run() { echo "== $*"; env "$@" TCS_DEBUG=1 python scripts/run_c4.py 100 100 2>&1 | grep -E "spmm ms|tiled" | head -2; }
run TCS_TILE_R=256 TCS_TILE_SMEM_KB=110
run TCS_TILE_R=256 TCS_TILE_SMEM_KB=110 TCS_NO_INTERLEAVE=1
run TCS_TILE_R=256 TCS_TILE_SMEM_KB=110 TCS_TILE_H=128
run TCS_TILE_R=256 TCS_TILE_SMEM_KB=74
run TCS_TILE_R=128 TCS_TILE_SMEM_KB=55
run TCS_TILE_R=384 TCS_TILE_SMEM_KB=110
