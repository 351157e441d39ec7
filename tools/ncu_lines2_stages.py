"""Line ranges of csrc/arm_kernel.cuh per engine stage (shared by tools/ncu_lines2.py and tools/ncu_sass_lines.py; update when the file moves)."""
STAGES = [(97, 152, "warp helpers (shfl, warp_sum, sincos, rsqrt)"), (153, 221, "vector helpers (cross, mat_vec, quat, spi_mul)"),
          (222, 277, "cylinder separating-axis tests"), (278, 346, "convex_pair_dev (MPR call site)"),
          (347, 416, "kinematics"), (417, 557, "dynamics(CRB/RNE/act)"), (558, 840, "collide"), (841, 987, "make_rows"), (988, 1020, "contact_eval"),
          (1021, 1038, "contact_apply"), (1039, 1052, "jt_force"), (1053, 1089, "chol6/solve6"), (1090, 1219, "block_direction"), (1220, 1253, "dense_direction"),
          (1254, 1438, "solve(body)"), (1439, 1461, "touch"), (1462, 1521, "integrate"), (1522, 1565, "convex_fixup"), (1566, 1638, "substep glue"),
          (1639, 1848, "ur_ik"), (1849, 2600, "task/io")]
