"""Empty stand-in: only the Fortran frontend of the reference imports fparser."""
