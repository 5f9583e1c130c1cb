"""CPU oracle for the CONAN contact path -- test infrastructure only (see conan_oracle.c)."""
