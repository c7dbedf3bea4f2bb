# Same import paths as the reference's models/ directory.
