"""Install ``deepspeed_b200`` (and the ``deepspeed`` import-compatibility package).

    pip install .                      # builds the sm_100a kernel library and the host runtime, then installs
    DS_SKIP_BUILD_OPS=1 pip install .  # install the Python only; the libraries are built on first use

Reference: ``setup.py`` (``DS_BUILD_OPS`` pre-compilation, console scripts ``deepspeed`` / ``ds`` / ``ds_report`` ...). The
native libraries are plain shared objects loaded with ctypes, so the build step is a call into ``ops/builder.py`` (nvcc
``-gencode arch=compute_100a,code=sm_100a``; g++ for the host runtime) rather than a torch C++ extension."""
import os
import re
import sys

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))


def _version():
    text = open(os.path.join(ROOT, "deepspeed_b200", "version.py")).read()
    return re.search(r'__version__\s*=\s*"([^"]+)"', text).group(1)


class BuildWithKernels(build_py):

    def run(self):
        if os.environ.get("DS_SKIP_BUILD_OPS", "0") != "1":
            sys.path.insert(0, ROOT)
            try:
                from deepspeed_b200.ops.builder import build_all
                build_all(verbose=True)
            except Exception as e:  # noqa: BLE001 - no nvcc on the build host: fall back to building on first use
                print(f"warning: native libraries were not pre-built ({e}); they are built on first use", file=sys.stderr)
        super().run()


setup(
    name="deepspeed_b200",
    version=_version(),
    description="B200-native (sm_100a, NVLink 5 / NVSwitch) training and inference framework with the DeepSpeed API",
    long_description=open(os.path.join(ROOT, "README.md")).read(),
    long_description_content_type="text/markdown",
    packages=find_packages(include=["deepspeed_b200", "deepspeed_b200.*", "deepspeed"]),
    package_data={"deepspeed_b200": ["ops/lib/*.so", "checkpoint/zero_to_fp32.py"]},
    data_files=[("share/deepspeed_b200/csrc", [os.path.join("csrc", f) for f in sorted(os.listdir("csrc"))
                                                if f.endswith((".cu", ".cuh", ".h"))]),
                ("share/deepspeed_b200/csrc/cpu", [os.path.join("csrc", "cpu", f)
                                                    for f in sorted(os.listdir(os.path.join("csrc", "cpu")))]),
                ("share/deepspeed_b200/include", [os.path.join("include", "dsb200.h")])],
    scripts=[os.path.join("bin", f) for f in ("deepspeed", "ds", "ds_bench", "ds_elastic", "ds_io", "ds_nvme_tune",
                                              "ds_report", "dsr", "ds_ssh")],
    python_requires=">=3.10",
    install_requires=["torch>=2.6", "numpy", "pydantic>=2", "packaging", "psutil", "hjson", "tqdm"],
    extras_require={"inf": ["transformers", "safetensors"], "dev": ["pytest", "pytest-timeout"]},
    cmdclass={"build_py": BuildWithKernels},
    classifiers=["Programming Language :: Python :: 3", "Environment :: GPU :: NVIDIA CUDA :: 12"],
)
