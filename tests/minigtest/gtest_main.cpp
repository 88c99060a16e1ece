// main() for the reference's test/*.cpp compiled against tests/minigtest (stands in for GTest::gtest_main).
#include <gtest/gtest.h>
int main(int argc, char** argv) {
    ::testing::InitGoogleTest(&argc, argv);
    return RUN_ALL_TESTS();
}
