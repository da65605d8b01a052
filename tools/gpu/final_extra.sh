python tests/devtools/time_voxel.py 2>&1 | tail -4
python tests/devtools/time_3d.py 2>&1 | tail -3 | cut -c1-200
