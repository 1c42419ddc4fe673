import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tools.timeline as tl
tl.run(100_000)
tl.run(100_000, mode=1, frames=12)
