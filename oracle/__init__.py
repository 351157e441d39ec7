"""CPU oracle -- TEST INFRASTRUCTURE ONLY (parity unpinned, see mjs_oracle.cpp)."""
