// fls_error.hpp — thread-local last-error slot behind flsgpu_last_error().
#pragma once
int fls_set_error(int status, const char* msg); // returns status
