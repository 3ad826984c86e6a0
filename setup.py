"""Packaging: the plugin is discovered by vLLM through the `vllm.platform_plugins` entry-point group
(reference setup.py:41-43).  The CUDA library is built in-tree by `python -m vllm_neuron_b200.build`."""
from setuptools import find_packages, setup

setup(
    name="vllm-neuron-b200",
    version="0.1.0",
    description="B200-native vLLM platform plugin with vllm-neuron's plugin surface (block-KV paged decode hot path)",
    packages=find_packages(include=["vllm_neuron_b200", "vllm_neuron_b200.*"]),
    package_data={"vllm_neuron_b200": ["libb200_paged.so", "csrc/*"]},
    python_requires=">=3.10",
    entry_points={"vllm.platform_plugins": ["b200 = vllm_neuron_b200:register"]},
)
