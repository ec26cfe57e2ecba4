// Test program for include/tinyfk_b200.hpp (the C++ view of the C ABI): drives the model the way the reference's own C++
// bench does (bench/bench_tinyfk.cpp:17-60: set_joint_angles, get_link_pose, get_jacobian) and prints every number as a hex
// float; tests/test_cpp_shim.py compares the output with the Python facade bit for bit.
#include <cstdio>
#include <fstream>
#include <sstream>

#include "tinyfk_b200.hpp"

using namespace tinyfk_b200;

static void dump(const char *name, const double *v, size_t n) {
  std::printf("%s", name);
  for (size_t i = 0; i < n; ++i) std::printf(" %a", v[i]);
  std::printf("\n");
}

int main(int argc, char **argv) {
  if (argc < 2) return 2;
  std::ifstream f(argv[1]);
  std::stringstream ss;
  ss << f.rdbuf();
  try {
    KinematicModel kin(ss.str());
    const std::vector<std::string> jn = {"r_shoulder_pan_joint", "r_shoulder_lift_joint", "r_upper_arm_roll_joint", "r_elbow_flex_joint",
                                         "r_forearm_roll_joint", "r_wrist_flex_joint",    "r_wrist_roll_joint"};
    const std::vector<size_t> jids = kin.get_joint_ids(jn);
    const std::vector<size_t> lids = kin.get_link_ids({"r_gripper_tool_frame", "r_elbow_flex_link", "base_link"});
    const size_t mylink = kin.add_new_link("mylink", lids[0], {0.1, 0.2, 0.3}, {0.3, 0.2, 0.1});
    if (argc > 2) { // no GPU needed up to here
      std::printf("ids %zu %zu %zu %zu\n", jids[0], lids[0], lids[2], mylink);
      return 0;
    }
    kin.set_joint_angles(jids, {0.564, 0.35, -0.74, -0.7, -0.7, -0.17, -0.63});
    kin.set_base_pose({0.1, 0.2, 0.05}, {0.02, -0.03, 0.3});
    for (size_t lid : {lids[0], lids[1], lids[2], mylink}) {
      Pose p;
      kin.get_link_pose(lid, p);
      dump("pose", p.position.data(), 3);
      dump("quat", p.rotation.data(), 4);
    }
    const Matrix J = kin.get_jacobian(lids[0], jids, RotationType::RPY, true);
    std::printf("jac %zu %zu\n", J.rows, J.cols);
    dump("J", J.data.data(), J.data.size());
    const Matrix J2 = kin.get_jacobian(mylink, jids, RotationType::IGNORE, false);
    dump("J2", J2.data.data(), J2.data.size());
    // a batch of three PLANER configurations
    std::vector<double> qs;
    for (int i = 0; i < 3; ++i)
      for (int c = 0; c < 10; ++c) qs.push_back(0.1 * (c + 1) - 0.37 * i);
    Matrix P, Jb;
    kin.solve_fk(qs, 3, {lids[0], mylink}, jids, RotationType::XYZW, BaseType::PLANER, true, P, Jb);
    dump("P", P.data.data(), P.data.size());
    dump("Jb", Jb.data.data(), Jb.data.size());
    try {
      kin.get_link_ids({"no_such_link"});
      return 3;
    } catch (const std::invalid_argument &) {
      std::printf("unknown link: invalid_argument\n");
    }
  } catch (const std::exception &e) {
    std::printf("error: %s\n", e.what());
    return 1;
  }
  return 0;
}
