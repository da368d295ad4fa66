# Round-1 run J: full GPU suite incl. the full-size C2 property test and the ported loader tests; smoke(); timing of the suite.
set -x
time python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r01j_pytest.log 2>&1; tail -25 gpurun_out/r01j_pytest.log
python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/r01j_smoke.log 2>&1; echo rc=$?; cat gpurun_out/r01j_smoke.log
