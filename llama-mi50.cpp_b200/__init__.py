"""ggml-b200: a B200-native (sm_100a) ggml backend for the quantized token-generation path.

The product is two shared libraries built in this directory by csrc/Makefile:

  libb200_kernels.so  hand-written CUDA kernels behind the C ABI of include/b200_kernels.h
  libggml-b200.so     the ggml backend plugin (ggml_backend_reg / device / buffer type / buffer /
                      backend vtables, include/b200_backend.h) that unmodified llama.cpp,
                      llama-bench and test-backend-ops load through GGML_BACKEND_PATH

This Python package is harness only: ctypes bindings of the C ABI for the parity tests
(`capi`), the synthetic GGUF writer (`gguf_synth`) and the llama-bench driver (`harness`).
There is no CPU fallback anywhere: loading fails loudly when the CUDA library is missing.

The directory name is not a valid Python identifier; import it with
`__graft_entry__.load_package()` (module name `llama_mi50_b200`).
"""
import os

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(PKG_DIR)
KERNELS_SO = os.environ.get("B200_KERNELS_SO") or os.path.join(PKG_DIR, "libb200_kernels.so")   # override: A/B runs of kernel builds (tools/)
BACKEND_SO = os.path.join(PKG_DIR, "libggml-b200.so")
REF_DIR = os.path.join(REPO_ROOT, "oracle", "_ref")   # unmodified reference host build (harness only)

__all__ = ["PKG_DIR", "REPO_ROOT", "KERNELS_SO", "BACKEND_SO", "REF_DIR"]
