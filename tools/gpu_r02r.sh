export PYTHONPATH=.
for ll in 1 0; do echo "PFMPE_LL=$ll"; PFMPE_LL=$ll python -c "
import tools.timeline as tl
tl.run(100_000)
tl.run(1_000_000, nout=200, frames=12)
"; done
