timeout 600 python -m pytest tests/test_lammps_plugin.py -x -q -m gpu 2>&1 | tail -3
python tools/lammps_water_bench.py --cells 46 46 46 --steps 200 --warm 50 --neighbor device --integrator device > gpurun_out/r2_lammps_stock_292k_device_md_final.log 2>&1
tail -11 gpurun_out/r2_lammps_stock_292k_device_md_final.log
python tools/lammps_water_bench.py --cells 46 46 46 --steps 200 --warm 50 --neighbor device --integrator device --weight-scale 0.25 > gpurun_out/r2_lammps_stock_292k_device_md_tame.log 2>&1
tail -11 gpurun_out/r2_lammps_stock_292k_device_md_tame.log
