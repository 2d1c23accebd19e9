"""Drop-in layer modules (surface S1 of SURVEY.md section 8b): the reference's class names, constructor
signatures, parameter names and forward signatures, with the arithmetic in libb200tts.so."""
