"""Import stub (test infrastructure): just enough of `diffusers` for the reference's two model
files to import in a container that does not have diffusers installed. No arithmetic lives here."""
class AutoencoderKL:  # imported by name only
    pass
