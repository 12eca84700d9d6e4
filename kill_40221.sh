ssh test@127.0.0.1 "kill -9 1920"
ssh test@127.0.0.1 "kill -9 1921"
ssh test@127.0.0.1 "kill -9 1922"
kill -9 1591
