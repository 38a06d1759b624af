// TEST INFRASTRUCTURE ONLY.  Stand-in for the reference's src/test/test_clore.h: kawpow_tests.cpp only names the
// BasicTestingSetup fixture (chain parameters, ECC start-up - nothing the KawPow tests touch); the real header pulls in
// the whole node.  See boost/test/unit_test.hpp beside this file.
#pragma once
struct BasicTestingSetup
{
};
