"""Empty stand-in: only the Octave frontend of the reference imports ply."""
