import sys

from .tools import main

sys.exit(main())
