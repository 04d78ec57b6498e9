"""Throw-away stand-in for the `deap` package (absent from this image).

TEST TOOLING ONLY: lets the *unmodified* reference front end
(`evostencils.grammar.multigrid.generate_primitive_set`,
`evostencils.code_generation.exastencils.ProgramGenerator.generate_cycle_function`)
be imported in this container so that golden fixtures can be generated from it.
Nothing under `evostencils_b200/` imports this.
"""
