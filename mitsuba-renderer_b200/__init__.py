"""mitsuba-renderer_b200 — CUDA (sm_100a) path-tracing hot path behind Mitsuba 0.2.1's
SampleIntegrator plugin API. Host-side Python mirror of the C ABI in include/mtsgpu.h."""
