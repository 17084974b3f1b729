python tools/diag_parity.py > gpurun_out/diag_step.txt 2>&1; WB200_NO_STEP_GRAPH=1 python tools/diag_parity.py > gpurun_out/diag_nostep.txt 2>&1; WB200_STEP_NO_GRAPH=1 WB200_NO_PDL=1 python tools/diag_parity.py > gpurun_out/diag_nopdl.txt 2>&1
cat gpurun_out/diag_step.txt gpurun_out/diag_nostep.txt gpurun_out/diag_nopdl.txt
