"""Packaging (role of /root/reference/setup.py:1-45): installs the package with its ``conf/``
tree as package data.  The CUDA library is built separately with
``python -m distributed_learning_simulator_b200.build`` (nvcc, sm_100a)."""
import setuptools

setuptools.setup(
    name="distributed_learning_simulator_b200",
    version="0.1.0",
    description="B200-native single-host simulator for distributed and federated learning",
    packages=setuptools.find_packages(include=["distributed_learning_simulator_b200*"]),
    package_data={"distributed_learning_simulator_b200": ["conf/*.yaml", "conf/*/*.yaml", "conf/*/*/*.yaml",
                                                         "csrc/*.cu", "csrc/*.cuh", "lib/*.so"]},
    include_package_data=True,
    python_requires=">=3.11",
    install_requires=["torch>=2.0", "pyyaml", "dill"],
)
