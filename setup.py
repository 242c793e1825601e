"""Packaging for torchgpipe_b200. ``build_ext`` / ``build_py`` first run ``make -C csrc``, which
compiles the sm_100a kernels into ``torchgpipe_b200/libtgb200.so`` (a plain C-ABI shared object
loaded through ctypes - no torch C++ extension, so no ABI coupling to a torch build)."""
import os
import subprocess

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

HERE = os.path.dirname(os.path.abspath(__file__))

about = {}
with open(os.path.join(HERE, 'torchgpipe_b200', '__version__.py')) as f:
    exec(f.read(), about)


class BuildWithKernels(build_py):
    def run(self):
        if os.environ.get('TGB_SKIP_NATIVE') != '1':
            subprocess.check_call(['make', '-C', os.path.join(HERE, 'csrc'), '-j8'])
        super().run()


setup(
    name='torchgpipe_b200',
    version=about['__version__'],
    description='GPipe for PyTorch, rebuilt for NVIDIA B200 (sm_100a)',
    license='BSD-3-Clause',
    packages=find_packages(include=['torchgpipe_b200', 'torchgpipe_b200.*']),
    package_data={'torchgpipe_b200': ['py.typed', 'libtgb200.so']},
    py_modules=['torchgpipe_balancing'],
    python_requires='>=3.9',
    install_requires=['torch>=2.4'],
    cmdclass={'build_py': BuildWithKernels},
    zip_safe=False,
)
