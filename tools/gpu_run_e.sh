mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -x -k "column_stacked or tap_stacked_equals" > gpurun_out/r02e_pytest_final.log 2>&1; echo "final tests exit $?"; tail -15 gpurun_out/r02e_pytest_final.log
timeout 300 python - <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, '.')
from argparse import Namespace
import isosurfacesuperresolution_b200 as isr
w = dict(np.load('tests/golden/enhancenet_weights_unshaded_seed1234.npz'))
for fin in (1, 3):
    isr._lib.set_tunable("final_stacked", fin)
    net = isr.createNetwork("EnhanceNet", 4, 101, [0,1,2,3,4], 6, Namespace(upsample="bilinear", reconType="residual", useBN=False))
    net.load_state_dict({k: torch.from_numpy(v) for k, v in w.items()}); net = net.cuda().eval()
    h = net._native()
    for shape in ((4, 512, 512), (16, 256, 256), (1, 270, 480)):
        print("final_stacked", fin, shape, "output conv ms", h.time_conv(*shape, "hybrid", scale=5, iters=20), flush=True)
PY
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/r02e_pytest.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/r02e_pytest.log
