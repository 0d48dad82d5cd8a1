// ORACLE / test infrastructure: stand-in so that the reference sources compile without ROS.
#pragma once
#include <assert.h>
#define ROS_ASSERT(cond) assert(cond)
#define ROS_ERROR(...) ((void)0)
#define ROS_INFO(...) ((void)0)
#define ROS_WARN(...) ((void)0)
