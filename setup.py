"""Packaging (reference: setup.py + requirements.txt).

The reference's setup.py only packages ``data/`` and compiles one Cython file; its modules use flat imports, so
the package works only from the checkout. Here ``hetseq_b200`` is a proper package; building it means compiling
the two native libraries in-tree (nvcc for sm_100a, g++), which ``build_py`` triggers -- the same code path as
``__graft_entry__.build()``. No Cython, no torch extension.

    pip install -e .            # or: python setup.py build_py
"""
import os
import sys

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

HERE = os.path.dirname(os.path.abspath(__file__))


class BuildWithNativeLibs(build_py):
    def run(self):
        sys.path.insert(0, HERE)
        try:
            from hetseq_b200.ops import build as native_build
            native_build.build_native(verbose=True)
            try:
                native_build.build_cuda(verbose=True)
            except RuntimeError as e:      # no nvcc: the CPU paths still work, say so loudly
                print('WARNING: CUDA kernel library not built ({}); GPU runs need it'.format(str(e).splitlines()[0]))
        finally:
            sys.path.pop(0)
        super().run()


setup(
    name='hetseq_b200',
    version='0.1.0',
    description='B200-native (sm_100a) synchronous data-parallel training framework with the capabilities of HetSeq',
    packages=find_packages(include=['hetseq_b200', 'hetseq_b200.*']),
    package_data={'hetseq_b200': ['lib/*.so', 'native/*.cpp']},
    python_requires='>=3.10',
    install_requires=['torch>=2.4', 'numpy'],
    extras_require={'hdf5': ['h5py'], 'mnist': ['torchvision'], 's3': ['boto3'], 'test': ['pytest']},
    entry_points={'console_scripts': ['hetseq-b200-train = hetseq_b200.train:cli_main',
                                      'hetseq-b200-eval-mnist = hetseq_b200.eval_mnist:main']},
    cmdclass={'build_py': BuildWithNativeLibs},
)
