// ORACLE — TEST INFRASTRUCTURE ONLY (see field.hpp header). Placeholder translation unit for
// the multi-threaded CPU baselines (filled in below by later commits).
