#!/bin/bash
# BASELINE's Astaroth target config per GPU; a wrapper because torchrun's own parser swallows "--star" (prefix of --start-method)
exec "$(dirname "$0")/../build/bin/astaroth-sim" --x ${AX:-256} --y ${AY:-256} --z ${AZ:-256} --fields 8 --double --star19 --iters ${ITERS:-20} ${EXTRA}
