"""Mirror of the reference's `pdvc/ops` package (functions/ and modules/)."""
