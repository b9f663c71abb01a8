#!/bin/bash
# end-of-session check: the driver's own sequence (gpu tests, smoke, bench both arms)
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/final_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/final_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep smoke
python bench.py --impl reference > gpurun_out/final_ref.json 2> gpurun_out/final_ref.err; echo "ref rc=$?"
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"; wc -l gpurun_out/final_bench.json gpurun_out/final_ref.json
