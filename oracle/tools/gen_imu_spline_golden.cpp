// Generates tests/golden/imu_spline.json: the synthetic IMU stream and ground-truth states that the reference's own IMU
// preintegration test builds (bs_models/tests/imu_preintegration_tests.cpp:77-160, class Data), with the reference's
// vendored generator: basalt::Se3Spline<5> (dependencies/basalt-headers-mirror), knot interval 10 s, 15 random knots,
// 20 s, IMU at 100 Hz sampled in the middle of every interval, gravity (0, 0, -9.80665).
// Test infrastructure: reads /root/reference at BUILD time only (this container); the JSON it writes is committed, so
// nothing on the GPU box needs the reference. Build + run: oracle/tools/gen_imu_spline_golden.sh
#include <basalt/spline/se3_spline.h>

#include <cstdio>

int main(int argc, char** argv) {
  const char* out = argc > 1 ? argv[1] : "imu_spline.json";
  const int num_knots = 15;
  const int64_t t0 = 0, knot_interval = 10000000000LL, duration = 20000000000LL, dt_ns = 10000000LL;
  const int64_t t_end = t0 + duration;
  const Eigen::Vector3d g(0, 0, -9.80665);   // bs_common GRAVITY_WORLD (utils.h:20-24)
  basalt::Se3Spline<5> spline(knot_interval, t0);
  spline.genRandomTrajectory(num_knots);
  FILE* f = std::fopen(out, "w");
  if (!f) return 1;
  std::fprintf(f, "{\n \"source\": \"basalt::Se3Spline<5>::genRandomTrajectory(15), knot interval 10 s; generator oracle/tools/gen_imu_spline_golden.cpp after bs_models/tests/imu_preintegration_tests.cpp:77-160\",\n");
  std::fprintf(f, " \"dt\": 0.01, \"gravity\": [0, 0, -9.80665],\n \"imu\": [\n");
  bool first = true;
  for (int64_t t = t0; t < t_end + dt_ns; t += dt_ns) {
    const int64_t tm = t + dt_ns / 2;   // state in the middle of the interval, stamped at its start (:93-104)
    const Sophus::SE3d pose = spline.pose(tm);
    const Eigen::Vector3d w = spline.rotVelBody(tm);
    const Eigen::Vector3d a = pose.so3().inverse() * (spline.transAccelWorld(tm) - g);
    std::fprintf(f, "%s  [%.17g, %.17g, %.17g, %.17g, %.17g, %.17g, %.17g]", first ? "" : ",\n", t * 1e-9, w[0], w[1], w[2], a[0], a[1], a[2]);
    first = false;
  }
  std::fprintf(f, "\n ],\n \"states\": [\n");
  const int64_t ts[3] = {t0, (t0 + t_end) / 2, t_end};   // IS1, IS2, IS3 (:82-84, :126-150)
  for (int k = 0; k < 3; ++k) {
    const Sophus::SE3d pose = spline.pose(ts[k]);
    const Eigen::Quaterniond q(pose.so3().matrix());
    const Eigen::Vector3d p = pose.translation(), v = spline.transVelWorld(ts[k]);
    std::fprintf(f, "  {\"t\": %.17g, \"q\": [%.17g, %.17g, %.17g, %.17g], \"p\": [%.17g, %.17g, %.17g], \"v\": [%.17g, %.17g, %.17g]}%s\n", ts[k] * 1e-9,
                 q.w(), q.x(), q.y(), q.z(), p[0], p[1], p[2], v[0], v[1], v[2], k < 2 ? "," : "");
  }
  std::fprintf(f, " ]\n}\n");
  std::fclose(f);
  return 0;
}
