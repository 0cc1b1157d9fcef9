# Sphinx configuration (reference: /root/reference/docs/conf.py). Build with
#   python -m sphinx -b html docs docs/_build
# Sphinx is not part of the offline image; the pages are plain reST + autodoc.
import os
import sys

sys.path.insert(0, os.path.abspath('..'))

project = 'kungfu-b200'
author = 'kungfu-b200 developers'
extensions = ['sphinx.ext.autodoc', 'sphinx.ext.napoleon', 'sphinx.ext.viewcode']
autodoc_mock_imports = ['torch']  # API pages build without a GPU stack
autodoc_member_order = 'bysource'
html_theme = 'alabaster'
exclude_patterns = ['_build']
master_doc = 'index'
