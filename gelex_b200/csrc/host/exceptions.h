// Exception types of the host engine.  Same names and hierarchy as the reference (include/gelex/exception.h:26-117)
// so that a caller that catches gelex::FileFormatException etc. keeps working when it switches namespaces, and so
// the CLI's single catch (apps/main.cpp:73-98: log the message, exit code 1) behaves identically.
#pragma once
#include <stdexcept>
#include <string>

#include "../../../include/ggrm.h"

namespace gelex::b200 {

class GelexException : public std::runtime_error {
   public:
    explicit GelexException(const std::string& m) : std::runtime_error(m) {}
};
#define GGRM_DEFINE_EXC(Name)                                          \
    class Name : public GelexException {                               \
       public:                                                         \
        explicit Name(const std::string& m) : GelexException(m) {}     \
    }
GGRM_DEFINE_EXC(FileNotFoundException);
GGRM_DEFINE_EXC(FileOpenException);
GGRM_DEFINE_EXC(FileWriteException);
GGRM_DEFINE_EXC(FileExistsException);
GGRM_DEFINE_EXC(FileFormatException);
GGRM_DEFINE_EXC(DataParseException);
GGRM_DEFINE_EXC(ArgumentValidationException);
GGRM_DEFINE_EXC(InvalidInputException);
GGRM_DEFINE_EXC(ColumnRangeException);
GGRM_DEFINE_EXC(InvalidOperationException);
GGRM_DEFINE_EXC(InconsistentColumnCountException);
#undef GGRM_DEFINE_EXC

// C-ABI status -> exception (include/ggrm.h ggrm_status documents the mapping).
[[noreturn]] inline void throw_status(int status, const std::string& msg) {
    switch (status) {
        case GGRM_ERR_INVALID_ARGUMENT: throw InvalidInputException(msg);
        case GGRM_ERR_COLUMN_RANGE: throw ColumnRangeException(msg);
        case GGRM_ERR_INVALID_OPERATION: throw InvalidOperationException(msg);
        case GGRM_ERR_OUT_OF_MEMORY: throw std::bad_alloc();
        case GGRM_ERR_FILE_FORMAT: throw FileFormatException(msg);
        default: throw GelexException(msg);
    }
}

}  // namespace gelex::b200
