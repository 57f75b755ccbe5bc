#!/bin/bash
# Builds the fixture generator against the reference's vendored basalt / Eigen / Sophus headers and writes
# tests/golden/imu_spline.json. Run in the build container (needs /root/reference); outputs go to oracle/_ref/.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"; ROOT="$HERE/../.."
B=/root/reference/dependencies/basalt-headers-mirror
mkdir -p "$ROOT/oracle/_ref"
g++ -std=c++17 -O2 -I$B/include -I$B/thirdparty/eigen -I$B/thirdparty/Sophus -I$B/thirdparty/cereal/include \
    -o "$ROOT/oracle/_ref/gen_imu_spline_golden" "$HERE/gen_imu_spline_golden.cpp"
"$ROOT/oracle/_ref/gen_imu_spline_golden" "$ROOT/tests/golden/imu_spline.json"
