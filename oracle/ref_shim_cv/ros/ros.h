// ORACLE / test infrastructure: stand-in so that the reference sources compile without ROS (nothing of it is used on the extractor path).
#pragma once
