"""nano-vllm-go B200 backend: host-side mirror of the reference's ModelRunner
interface over the C-ABI library libnvllm_b200.so (hand-written sm_100a CUDA).

The directory name carries a hyphen (it is the name the build contract asks
for), so it is imported through ``nvllm_pkg.py`` at the repo root, which
registers it as the module ``nano_vllm_go_b200``.
"""
__all__ = ["config", "weights", "lib", "runner", "build"]
