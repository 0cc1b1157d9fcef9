"""pip install . : builds the native runtime and kernels in-tree, installs the
``kungfu_b200`` package (plus the ``kungfu`` alias) and the ``kungfu-run`` /
``kungfu-config-server`` console scripts.
Reference packaging: /root/reference/setup_tensorflow.py (CMake + Go),
/root/reference/setup_pytorch.py:10-53; no CMake, Go or TensorFlow here."""
import os
import sys

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))


class BuildNative(build_py):
    def run(self):
        sys.path.insert(0, ROOT)
        from kungfu_b200 import _build
        _build.build_host()
        if os.environ.get('KUNGFU_BUILD_CUDA', '1') != '0':
            _build.build_cuda()
        super().run()


setup(
    name='kungfu_b200',
    version='0.1.0',
    description='Blackwell-native adaptive data-parallel training (KungFu API) for 8xB200',
    packages=find_packages(include=['kungfu_b200', 'kungfu_b200.*', 'kungfu', 'kungfu.*']),
    package_data={'kungfu_b200': ['lib/*.so', 'bin/*']},
    cmdclass={'build_py': BuildNative},
    entry_points={'console_scripts': ['kungfu-run=kungfu_b200.cmd:run',
                                      'kungfu-config-server=kungfu_b200.cmd:config_server']},
    python_requires='>=3.8',
)
