ssh test@127.0.0.1 "kill -9 2336"
ssh test@127.0.0.1 "kill -9 2337"
kill -9 2109
