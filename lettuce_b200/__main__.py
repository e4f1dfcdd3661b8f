"""`python -m lettuce_b200 ...` == the reference's `lettuce ...` console script (setup.py:28-32)"""
import sys

from .cli import main

sys.exit(main())
