from p2pfl_b200.cli import main

main()
