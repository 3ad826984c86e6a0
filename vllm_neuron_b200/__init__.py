"""B200-native vLLM platform plugin with vllm-neuron's plugin surface.

Entry point (setup.py): ``vllm.platform_plugins: b200 = vllm_neuron_b200:register``.
Mirrors vllm_neuron/__init__.py:8-23 of the reference: a cheap device probe that returns the
platform's qualified name, or ``None`` plus a UserWarning when the device is absent.  vLLM calls
this in every process, so it must stay import-light: no torch, no CUDA context.
"""
import glob
import warnings

__version__ = "0.1.0"
PLATFORM_QUALNAME = "vllm_neuron_b200.platform.B200Platform"


def _sm100_present() -> bool:
    """True iff an NVIDIA device of compute capability 10.x is visible (NVML; no CUDA context)."""
    if not glob.glob("/dev/nvidia[0-9]*"):
        return False
    try:
        import pynvml

        pynvml.nvmlInit()
        try:
            for i in range(pynvml.nvmlDeviceGetCount()):
                major, _ = pynvml.nvmlDeviceGetCudaComputeCapability(pynvml.nvmlDeviceGetHandleByIndex(i))
                if major == 10:
                    return True
            return False
        finally:
            pynvml.nvmlShutdown()
    except Exception:
        # NVML unavailable: the device nodes exist, let the platform's own check decide later
        return True


def register():
    """Return the platform qualname iff a B200-class device is present, else warn and return None."""
    if not _sm100_present():
        warnings.warn(
            "No sm_100 (B200) devices found. Skipping B200 plugin registration.",
            category=UserWarning,
        )
        return None
    return PLATFORM_QUALNAME
