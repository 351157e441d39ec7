"""Line ranges of csrc/arm_kernel.cuh per engine stage (shared by tools/ncu_lines2.py and tools/ncu_sass_lines.py; update when the file moves)."""
STAGES = [(101, 166, "warp helpers (shfl, warp_sum, sincos, rsqrt)"), (167, 235, "vector helpers (cross, mat_vec, quat, spi_mul)"),
          (236, 291, "cylinder separating-axis tests"), (292, 364, "convex_pair_dev (MPR call site)"),
          (365, 434, "kinematics"), (435, 642, "dynamics(CRB/RNE/act)"), (643, 925, "collide"), (926, 1108, "make_rows"), (1109, 1141, "contact_eval"),
          (1142, 1159, "contact_apply"), (1160, 1173, "jt_force"), (1174, 1212, "chol6/solve6"), (1213, 1347, "block_direction"), (1348, 1381, "dense_direction"),
          (1382, 1598, "solve(body)"), (1599, 1621, "touch"), (1622, 1739, "integrate"), (1740, 1817, "convex_fixup"), (1818, 1884, "substep glue"),
          (1885, 2102, "ur_ik"), (2103, 2704, "task/io")]
