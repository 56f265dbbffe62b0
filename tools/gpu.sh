#!/bin/bash
# local wrapper: always rebuild the library before shipping the tree to the GPU box
set -e
cd /root/repo
python styletts2_b200/build.py > /dev/null
exec /usr/local/graft/bin/gpurun "$@"
