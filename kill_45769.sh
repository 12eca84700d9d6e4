ssh test@127.0.0.1 "kill -9 1998"
ssh test@127.0.0.1 "kill -9 1999"
ssh test@127.0.0.1 "kill -9 2000"
kill -9 1591
