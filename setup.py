"""``pip install -e .`` (optional): builds the native libraries in-tree with the same commands as
``__graft_entry__.build()`` and installs the ``hetu`` / ``hetuml`` / ``graphmix`` packages.
The reference builds with CMake (CMakeLists.txt, cmake/config.cmake); here nvcc / g++ are driven
directly because every library is a handful of translation units."""
import os
import sys

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))


class BuildNative(build_py):
    def run(self):
        sys.path.insert(0, ROOT)
        import __graft_entry__
        __graft_entry__.build()
        super().run()


setup(
    name='hetu-b200',
    version='0.1.0',
    description='Blackwell-native distributed deep-learning framework with the capabilities of Hetu',
    packages=find_packages(include=['hetu', 'hetu.*', 'hetuml', 'hetuml.*', 'graphmix', 'graphmix.*', 'graphboard', 'graphboard.*',
                                    'hetu_61a7_b200', 'hetu_61a7_b200.*']),
    package_data={'hetu_61a7_b200': ['lib/*.so']},
    scripts=['bin/heturun'],
    python_requires='>=3.9',
    install_requires=['numpy', 'torch', 'pyyaml', 'scipy'],
    cmdclass={'build_py': BuildNative},
)
