// internal.h -- shared between the translation units of libhmb200.so (not part of the ABI)
#pragma once
void hmb200i_set_error(const char *msg);
