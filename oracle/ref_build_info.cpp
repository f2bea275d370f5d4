// TEST INFRASTRUCTURE. The reference generates these four constants at CMake configure
// time (common/build-info.cpp.in); oracle/Makefile does not run CMake, so they are
// supplied here. Nothing else of the reference's `common` library is replaced.
int LLAMA_BUILD_NUMBER = 0;
char const *LLAMA_COMMIT = "reference";
char const *LLAMA_COMPILER = "gcc";
char const *LLAMA_BUILD_TARGET = "x86_64-linux-gnu";
