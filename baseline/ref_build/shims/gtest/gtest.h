// shim: driver/simple_id_mapper.cpp includes gtest in production code without using it
#ifndef MINIPS_SHIM_GTEST_H_
#define MINIPS_SHIM_GTEST_H_
#define FRIEND_TEST(a, b)
#endif
