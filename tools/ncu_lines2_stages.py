"""Line ranges of csrc/arm_kernel.cuh per engine stage (shared by tools/ncu_lines2.py and tools/ncu_sass_lines.py; update when the file moves)."""
STAGES = [(101, 160, "warp helpers (shfl, warp_sum, sincos, rsqrt)"), (161, 231, "vector helpers (cross, mat_vec, quat, spi_mul)"),
          (232, 287, "cylinder separating-axis tests"), (288, 360, "convex_pair_dev (MPR call site)"),
          (361, 430, "kinematics"), (431, 638, "dynamics(CRB/RNE/act)"), (639, 921, "collide"), (922, 1104, "make_rows"), (1105, 1137, "contact_eval"),
          (1138, 1155, "contact_apply"), (1156, 1169, "jt_force"), (1170, 1208, "chol6/solve6"), (1209, 1343, "block_direction"), (1344, 1377, "dense_direction"),
          (1378, 1594, "solve(body)"), (1595, 1617, "touch"), (1618, 1735, "integrate"), (1736, 1813, "convex_fixup"), (1814, 1880, "substep glue"),
          (1881, 2098, "ur_ik"), (2099, 2700, "task/io")]
