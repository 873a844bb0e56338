"""oracle — TEST INFRASTRUCTURE.  CPU checkers for the ORB frontend hot path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package; the product (orb_slam3_ros2_b200) never does.
"""
