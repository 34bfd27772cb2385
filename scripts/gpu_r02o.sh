#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
timeout 900 python scripts/gpu_icfp_diag2.py chess:f64 dice:f32 dice:f64 chess:f32 2>&1 | cut -c1-1500
