"""Builds libsepb_host.so: the C++ service layer (separator_services.cpp) + the ctypes test shim, linked against the
in-tree libsepb200.so.  ros_node.cpp is only compiled where roscpp exists (-DSEPB_WITH_ROS in a catkin build)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
PKG = os.path.join(ROOT, "multi_robot_slam_separators_b200")
SO = os.path.join(HERE, "libsepb_host.so")


def build(force=False):
    srcs = [os.path.join(HERE, f) for f in ("separator_services.cpp", "c_shim.cpp", "ros_node.cpp")]
    deps = srcs + [os.path.join(HERE, f) for f in ("separator_services.hpp", "sepb_msgs.hpp")] + \
        [os.path.join(ROOT, "include", "sepb200.h")]
    if not force and os.path.exists(SO) and all(os.path.getmtime(d) <= os.path.getmtime(SO) for d in deps):
        return SO
    cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-fPIC", "-shared"] + srcs + \
          ["-L" + PKG, "-lsepb200", "-Wl,-rpath," + PKG, "-o", SO]
    subprocess.check_call(cmd)
    return SO


if __name__ == "__main__":
    print(build(force=True))
