"""vda_b200: B200 (sm_100a) kernels + host mirror of Video Depth Anything's per-window inference path."""
__all__ = ["lib", "build"]
