import sys
from .b200run import main
sys.exit(main())
