ssh test@127.0.0.1 "kill -9 1780"
ssh test@127.0.0.1 "kill -9 1781"
ssh test@127.0.0.1 "kill -9 1782"
kill -9 1591
