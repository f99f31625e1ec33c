/* TEST INFRASTRUCTURE — stands in for cslam/config.h (ROS parameters) in the oracle build of ORBmatcher.cpp. */
#pragma once
#include <iostream>
#include <mutex>
#include <cslam/estd.h>
