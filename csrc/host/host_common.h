#pragma once
namespace idcbh {
void set_error(const char* fmt, ...);
}
