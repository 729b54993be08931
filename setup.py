# -*- coding: utf-8 -*-
"""Setup script for distributed-dot-product (B200-native).

``python setup.py build_ext --inplace`` (or ``pip install -e .``) compiles the CUDA sources for sm_100a with nvcc
through build.py and places ``libddp_b200.so`` inside the package.  Unlike the reference's setup.py
(reference: setup.py:37-64) every sub-package is listed (``utils`` has an ``__init__``), nothing is fetched at
build time (no ``setup_requires``), and there are no stale winpty/backports entries.
"""
import ast
import os
import sys

from setuptools import Command, find_packages, setup
from setuptools.command.build_py import build_py

HERE = os.path.abspath(os.path.dirname(__file__))


def get_version(module='distributed_dot_product'):
    """Reads VERSION_INFO without importing the package."""
    with open(os.path.join(HERE, module, '__init__.py'), 'r') as f:
        for line in f.read().split('\n'):
            if line.startswith('VERSION_INFO'):
                return '.'.join(map(str, ast.literal_eval(line.split('=')[-1].strip())))
    raise RuntimeError('VERSION_INFO not found')


def _build_native():
    sys.path.insert(0, HERE)
    import build as _build
    _build.build_lib()


class BuildNative(Command):
    description = 'compile csrc/ for sm_100a with nvcc (build.py)'
    user_options = [('inplace', 'i', 'ignored: the library is always built in-tree')]

    def initialize_options(self):
        self.inplace = 1

    def finalize_options(self):
        pass

    def run(self):
        _build_native()


class BuildPyWithNative(build_py):
    def run(self):
        _build_native()
        build_py.run(self)


setup(
    name='distributed-dot-product-b200',
    version=get_version(),
    keywords=['transformer', 'attention', 'sequence parallel', 'blackwell', 'pytorch'],
    license='MIT',
    description='Sequence-parallel dot-product attention operators with fused collective tcgen05 kernels for B200',
    long_description=open(os.path.join(HERE, 'README.md')).read(),
    long_description_content_type='text/markdown',
    packages=find_packages(exclude=['tests*', 'baseline*', 'benchmarks*', 'tools*']),
    package_data={'distributed_dot_product': ['_lib/*.so']},
    include_package_data=True,
    install_requires=['torch'],
    cmdclass={'build_ext': BuildNative, 'build_native': BuildNative, 'build_py': BuildPyWithNative},
    classifiers=[
        'Development Status :: 4 - Beta',
        'License :: OSI Approved :: MIT License',
        'Intended Audience :: Science/Research',
        'Operating System :: POSIX :: Linux',
        'Programming Language :: Python :: 3.12',
        'Environment :: GPU :: NVIDIA CUDA',
    ],
)
